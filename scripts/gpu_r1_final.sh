# end-of-round evidence: gpu tests, smoke, reference arm, default bench, launch list, full captures, SW sweep
set -x
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; tail -2 gpurun_out/pytest_gpu.log
python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; tail -1 gpurun_out/smoke.log
timeout 1200 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_mem_ref.json 2> gpurun_out/bench_mem_ref.err; echo "ref rc=$?"
timeout 1200 python bench.py > gpurun_out/bench_mem.json 2> gpurun_out/bench_mem.err; echo "mem rc=$?"
tail -c 600 gpurun_out/bench_mem.json
for kb in "16 100" "16 50" "16 16" "8 100" "8 16"; do set -- $kb
  timeout 600 python bench.py --workload bsw --kind $1 --band $2 --pairs 10000000 --steps 3 --warmup 3 --no-cpu > gpurun_out/bench_bsw_k$1_w$2.json 2> gpurun_out/bench_bsw_k$1_w$2.err; echo "bsw $1 $2 rc=$?"
done
timeout 900 python bench.py --workload bsw --kind 16 --band 100 --steps 3 --warmup 3 > gpurun_out/bench_bsw.json 2> gpurun_out/bench_bsw.err
CMD="python bench.py --steps 2 --warmup 1 --no-cpu"
ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/launches_mem.csv $CMD > gpurun_out/ncu_launch_mem.log 2>&1; echo "ncu list rc=$?"
CMD="python bench.py --steps 1 --warmup 1 --no-cpu"
for K in bsw_fast_kernel seed_collect_kernel chain_build_kernel; do
  ncu --set full --clock-control none --import-source on -k regex:$K -c 1 -f -o gpurun_out/prof_final_$K $CMD > gpurun_out/ncu_final_$K.log 2>&1; echo "$K rc=$?"
done
