# ERT evidence: gpu tests, live bench-scale parity against the reference, bench --ert, launch list, full capture
set -x
mkdir -p gpurun_out
python -m pytest tests/test_ert_gpu.py -m gpu -x -q > gpurun_out/pytest_ert_gpu.log 2>&1; tail -3 gpurun_out/pytest_ert_gpu.log
timeout 900 python scripts/ert_box_check.py 200000 50000 > gpurun_out/ert_box_check.json 2> gpurun_out/ert_box_check.err; echo "box check rc=$?"; tail -c 700 gpurun_out/ert_box_check.json; tail -3 gpurun_out/ert_box_check.err
timeout 1200 python bench.py --ert 1 --steps 5 --warmup 3 > gpurun_out/bench_mem_ert.json 2> gpurun_out/bench_mem_ert.err; echo "ert bench rc=$?"
tail -c 2500 gpurun_out/bench_mem_ert.json; tail -3 gpurun_out/bench_mem_ert.err
CMD="python bench.py --ert 1 --steps 2 --warmup 1 --no-cpu"
ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/launches_mem_ert.csv $CMD > gpurun_out/ncu_launch_mem_ert.log 2>&1; echo "ncu list rc=$?"
CMD="python bench.py --ert 1 --steps 1 --warmup 1 --no-cpu"
ncu --set full --clock-control none --import-source on -k regex:ert_seed_kernel -c 1 -f -o gpurun_out/prof_ert_seed $CMD > gpurun_out/ncu_ert_seed.log 2>&1; echo "ncu full rc=$?"
