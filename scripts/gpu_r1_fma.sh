set -x
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
timeout 900 python bench.py --workload seed --fma 0 --steps 3 --warmup 2 --no-cpu > gpurun_out/seed_nofma.json 2> gpurun_out/seed_nofma.err; echo "rc=$?"
timeout 900 python bench.py --workload seed --fma 1 --steps 3 --warmup 2 --no-cpu > gpurun_out/seed_fma.json 2> gpurun_out/seed_fma.err; echo "rc=$?"
cat gpurun_out/seed_nofma.json gpurun_out/seed_fma.json | cut -c1-1500
