# A/B run of differently-tuned builds of the seeding kernel (BWAGPU_LIB selects the build)
mkdir -p gpurun_out
for v in "" $VARIANTS; do
  BWAGPU_LIB=$PWD/bwa_mem_scale_b200/libbwagpu$v.so python bench.py --workload seed --fma 1 --steps 3 --warmup 2 --no-cpu 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('VAR [$v]', round(d['value']/1e6,2), 'Mreads/s kernel_ms', round(d['roofline']['kernel_ms'],2), 'frac', round(d['roofline']['frac'],3), 'ext/read', round(d['roofline']['bext_per_read'],1))"
done 2>&1 | tee gpurun_out/seed_ab.log
