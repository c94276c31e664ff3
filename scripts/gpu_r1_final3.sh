# final evidence of round 1: full gpu suite, smoke, default bench + reference arm, launch list of the default step,
# CIGAR bench + ncu summary of its kernels
set -x
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; tail -2 gpurun_out/pytest_gpu.log
python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; tail -2 gpurun_out/smoke.log
timeout 1200 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_mem_ref.json 2> gpurun_out/bench_mem_ref.err; echo "ref rc=$?"
timeout 1200 python bench.py > gpurun_out/bench_mem.json 2> gpurun_out/bench_mem.err; echo "mem rc=$?"
tail -c 500 gpurun_out/bench_mem.json
CMD="python bench.py --steps 2 --warmup 1 --no-cpu"
ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/launches_mem_final3.csv $CMD > gpurun_out/ncu_launch_mem_final3.log 2>&1; echo "ncu list rc=$?"
CMD="python bench.py --workload cigar --reads 300000 --steps 1 --warmup 1 --no-cpu"
ncu --set full --clock-control none --import-source on -k regex:cigar_kernel -c 4 -f -o gpurun_out/prof_cigar $CMD > gpurun_out/ncu_cigar.log 2>&1; echo "ncu cigar rc=$?"
