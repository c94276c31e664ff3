# A/B run of differently-tuned builds of the seeding kernel (BWAGPU_LIB selects the build)
for v in a b c d e f g h; do
  BWAGPU_LIB=$PWD/bwa_mem_scale_b200/libbwagpu_$v.so python bench.py --workload seed --steps 3 --warmup 2 --no-cpu 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('VAR $v', round(d['value']/1e6,2), 'Mreads/s kernel_ms', round(d['roofline']['kernel_ms'],2), 'frac', round(d['roofline']['frac'],3))"
done
