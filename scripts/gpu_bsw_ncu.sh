mkdir -p gpurun_out
CMD="python bench.py --workload bsw --steps 1 --warmup 1 --no-cpu"
$CMD > gpurun_out/plain_bsw.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:bsw_fast -c 1 -f -o gpurun_out/prof_bsw_fast $CMD > gpurun_out/ncu_bsw.log 2>&1
tail -2 gpurun_out/ncu_bsw.log | cut -c1-200
