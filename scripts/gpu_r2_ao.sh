# round 2, final launch list of one config-3 step (final code)
set -x
mkdir -p gpurun_out
CMD="python bench.py --reads 2500000 --chunk 2500000 --contexts 1 --steps 1 --warmup 1 --no-cpu"
timeout 500 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'^(?!fmb_|emfb_|eb_|fma_build|fm_layout).*' --csv --log-file gpurun_out/r2ao_launches.csv $CMD > gpurun_out/r2ao_launch.log 2>&1; echo "list rc=$?"
wc -l gpurun_out/r2ao_launches.csv
