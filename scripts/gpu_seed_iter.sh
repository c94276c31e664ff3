# quick iteration loop for the seeding kernels: parity, then the bench at two sizes
mkdir -p gpurun_out
python -m pytest tests/test_seed_gpu.py -m gpu -x -q 2>&1 | tail -3
python bench.py --workload seed --ref-mbp 20 --reads 200000 --steps 3 --warmup 2 --no-cpu 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('SMALL', d['value'], d['roofline'])"
python bench.py --workload seed --steps 3 --warmup 2 --no-cpu 2>&1 | tail -1 | tee gpurun_out/seed_iter.json | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('FULL', d['value'], d['e2e'], d['roofline'])"
