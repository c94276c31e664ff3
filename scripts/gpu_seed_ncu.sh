mkdir -p gpurun_out
CMD="python bench.py --workload seed --ref-mbp 100 --reads 1000000 --steps 1 --warmup 1 --no-cpu"
$CMD > gpurun_out/plain_seed.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:seed_collect -c 1 -f -o gpurun_out/prof_seed_collect $CMD > gpurun_out/ncu_full.log 2>&1
tail -2 gpurun_out/ncu_full.log | cut -c1-300
