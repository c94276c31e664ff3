# round 2, call Y: ncu --set full of the two top kernels (current code) + launch list of one step (index build skipped)
set -x
mkdir -p gpurun_out
CMD="python bench.py --reads 2500000 --chunk 2500000 --contexts 1 --steps 1 --warmup 1 --no-cpu"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:seed_collect_kernel -s 1 -c 1 -f -o gpurun_out/r2y_seed_collect_kernel $CMD > gpurun_out/r2y_ncu_seed.log 2>&1; echo "seed rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:bsw_fast_kernel -s 12 -c 1 -f -o gpurun_out/r2y_bsw_fast_kernel $CMD > gpurun_out/r2y_ncu_bsw.log 2>&1; echo "bsw rc=$?"
# launch list: skip the FM / EMF builders (first launches), list the kernels of the warm-up and timed steps
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'^(?!fmb_|emfb_|eb_|fma_build|fm_layout).*' --csv --log-file gpurun_out/r2y_launches.csv $CMD > gpurun_out/r2y_launch.log 2>&1; echo "list rc=$?"
wc -l gpurun_out/r2y_launches.csv
