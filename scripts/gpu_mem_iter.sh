mkdir -p gpurun_out
python -m pytest tests/test_bsw_gpu.py tests/test_c2a_gpu.py tests/test_chain_gpu.py -m gpu -x -q 2>&1 | tail -2
for env in "X=1" $EXTRA_ENVS; do
env $env python bench.py --steps 3 --warmup 2 --no-cpu 2>&1 | tail -1 | python -c "
import sys,json; d=json.loads(sys.stdin.read()); r=d.get('roofline',{})
print('$env', round(d['value']/1e6,2), 'Mreads/s', round(d['ms_per_step'],1), 'ms  e2e', round(d['e2e']['value']/1e6,2), {k:round(v,1) for k,v in r.get('kernel_ms_per_step',{}).items()})"
done 2>&1 | tee gpurun_out/mem_iter.log
