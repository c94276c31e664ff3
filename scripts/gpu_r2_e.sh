# round 2, call E: ncu captures of the three big kernels on the config-3 index (2 M reads, one chunk call each)
set -x
mkdir -p gpurun_out
CMD="python bench.py --reads 2000000 --steps 1 --warmup 1 --no-cpu"
for K in chain_build_kernel seed_collect_kernel; do
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:$K -s 2 -c 1 -f -o gpurun_out/r2e_$K $CMD > gpurun_out/r2e_ncu_$K.log 2>&1; echo "$K rc=$?"
done
# the SW kernel: the longest tier of the first chunk (launch 0 of the 16-bit kernel)
timeout 900 ncu --set full --clock-control none --import-source on -k regex:bsw_fast_kernel -s 0 -c 1 -f -o gpurun_out/r2e_bsw_fast_kernel $CMD > gpurun_out/r2e_ncu_bsw.log 2>&1; echo "bsw rc=$?"
ls -la gpurun_out/*.ncu-rep
