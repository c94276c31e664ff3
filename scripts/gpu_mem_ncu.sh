# ncu --set full captures of the three dominant kernels of the default (seed+extend) bench step
mkdir -p gpurun_out
CMD="python bench.py --steps 1 --warmup 1 --no-cpu"
$CMD > gpurun_out/plain_mem.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/plain_mem.log; exit 1; }
for K in bsw_fast_kernel seed_collect_kernel chain_build_kernel c2a_fill_kernel c2a_filter_kernel; do
  ncu --set full --clock-control none --import-source on -k regex:$K -c 1 -f -o gpurun_out/prof_mem_$K $CMD > gpurun_out/ncu_mem_$K.log 2>&1
  echo "$K rc=$?"
done
