# end-of-round evidence after the ERT work: gpu tests, smoke, reference arm, default bench, ERT bench, ERT captures
set -x
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; tail -2 gpurun_out/pytest_gpu.log
python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; tail -2 gpurun_out/smoke.log
timeout 1200 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_mem_ref.json 2> gpurun_out/bench_mem_ref.err; echo "ref rc=$?"
timeout 1200 python bench.py > gpurun_out/bench_mem.json 2> gpurun_out/bench_mem.err; echo "mem rc=$?"
tail -c 700 gpurun_out/bench_mem.json
timeout 1200 python bench.py --ert 1 --steps 5 --warmup 3 > gpurun_out/bench_mem_ert_v2.json 2> gpurun_out/bench_mem_ert_v2.err; echo "ert rc=$?"
timeout 1200 python bench.py --ert 1 --impl reference --steps 2 --warmup 1 > gpurun_out/bench_mem_ert_ref.json 2> gpurun_out/bench_mem_ert_ref.err; echo "ert ref rc=$?"
CMD="python bench.py --ert 1 --steps 1 --warmup 1 --no-cpu"
ncu --set full --clock-control none --import-source on -k regex:ert_seed_kernel -c 1 -f -o gpurun_out/prof_ert_seed_v2 $CMD > gpurun_out/ncu_ert_seed_v2.log 2>&1; echo "ncu full rc=$?"
