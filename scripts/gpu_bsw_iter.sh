mkdir -p gpurun_out
python -m pytest tests/test_bsw_gpu.py -m gpu -x -q 2>&1 | tail -3
for k in 16 8; do
python bench.py --workload bsw --kind $k --steps 3 --warmup 2 --no-cpu 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('BSW$k', round(d['value'],1), 'GCUPS  ms', round(d['ms_per_step'],2), 'e2e', round(d['e2e']['value'],1), d.get('roofline'))"
done
BWAGPU_BSW_PATH=thread python bench.py --workload bsw --steps 3 --warmup 2 --no-cpu 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('BSW16-v1', round(d['value'],1), 'GCUPS  ms', round(d['ms_per_step'],2))"
