set -x
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -3 gpurun_out/pytest_gpu.log
python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; tail -2 gpurun_out/smoke.log
timeout 1200 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_mem_ref.json 2> gpurun_out/bench_mem_ref.err; echo "ref rc=$?"
timeout 1200 python bench.py > gpurun_out/bench_mem.json 2> gpurun_out/bench_mem.err; echo "mem rc=$?"
tail -c 2500 gpurun_out/bench_mem.json; tail -5 gpurun_out/bench_mem.err
CMD="python bench.py --steps 2 --warmup 1 --no-cpu"
ncu --metrics gpu__time_duration.sum --clock-control none -c 2000 --csv --log-file gpurun_out/launches_mem.csv $CMD > gpurun_out/ncu_launch_mem.log 2>&1; echo "ncu rc=$?"
