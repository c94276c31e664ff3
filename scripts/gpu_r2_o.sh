set -x
mkdir -p gpurun_out
CMD="python bench.py --reads 2000000 --steps 1 --warmup 1 --no-cpu"
for K in chain_build_kernel c2a_filter_kernel c2a_filter_warp_kernel c2a_fill_kernel; do
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:^$K -s 2 -c 1 -f -o gpurun_out/r2o_$K $CMD > gpurun_out/r2o_ncu_$K.log 2>&1; echo "$K rc=$?"
done
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/r2o_launches.csv $CMD > gpurun_out/r2o_launch.log 2>&1; echo "list rc=$?"
