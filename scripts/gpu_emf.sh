mkdir -p gpurun_out
python -m pytest tests/test_emf_gpu.py -x -q > gpurun_out/pytest_emf.log 2>&1; tail -5 gpurun_out/pytest_emf.log
timeout 900 python bench.py --emf 1 --steps 3 --warmup 2 --no-cpu > gpurun_out/bench_mem_emf.json 2> gpurun_out/bench_mem_emf.err; echo "rc=$?"; tail -c 1800 gpurun_out/bench_mem_emf.json; tail -3 gpurun_out/bench_mem_emf.err
