set -x
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
timeout 900 python bench.py --workload seed > gpurun_out/bench_seed.json 2> gpurun_out/bench_seed.err; echo "seed rc=$?"
tail -c 3000 gpurun_out/bench_seed.json
timeout 600 python bench.py --workload bsw > gpurun_out/bench_bsw.json 2> gpurun_out/bench_bsw.err; echo "bsw rc=$?"
CMD="python bench.py --workload seed --ref-mbp 20 --reads 200000 --steps 1 --warmup 1 --no-cpu"
$CMD > gpurun_out/plain_seed_small.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches_seed.csv $CMD > gpurun_out/ncu_launch.log 2>&1
$CMD > gpurun_out/plain_seed_small2.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:seed_collect -c 1 -o gpurun_out/prof_seed_collect $CMD > gpurun_out/ncu_full.log 2>&1
ls -la gpurun_out
