# A/B of ert_seed_kernel builds (launch bounds / block size), bench --ert, kernel time per step from the profiling pass
mkdir -p gpurun_out
python -m pytest tests/test_ert_gpu.py -m gpu -x -q 2>&1 | tail -2
python __graft_entry__.py smoke 2>&1 | tail -2
for v in ""; do
  BWAGPU_LIB=$PWD/bwa_mem_scale_b200/libbwagpu$v.so python bench.py --ert 1 --steps 3 --warmup 3 --no-cpu > gpurun_out/bench_mem_ert_ab$v.json 2> gpurun_out/bench_mem_ert_ab$v.err
  python -c "
import json,sys; d=json.load(open('gpurun_out/bench_mem_ert_ab$v.json')); k=d['roofline']['kernel_ms_per_step']; print('variant[$v]', round(d['value']), round(d['ms_per_step'],1), 'ert_seed', round(k['ert_seed'],2), 'chain', round(k['chain_seeds'],1), 'bsw', round(k['bsw'],1))"
done
