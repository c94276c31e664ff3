# round 2, call D: config-3 workload, small trial then full scale
set -x
mkdir -p gpurun_out
BWAGPU_VERBOSE=1 timeout 600 python bench.py --ref-mbp 200 --reads 2000000 --steps 2 --warmup 1 > gpurun_out/r2d_small.json 2> gpurun_out/r2d_small.err; echo "small rc=$?"; tail -5 gpurun_out/r2d_small.err; cat gpurun_out/r2d_small.json | head -c 3000
BWAGPU_VERBOSE=1 timeout 1500 python bench.py --steps 2 --warmup 1 > gpurun_out/r2d_full.json 2> gpurun_out/r2d_full.err; echo "full rc=$?"; tail -8 gpurun_out/r2d_full.err; cat gpurun_out/r2d_full.json | head -c 4000
nvidia-smi --query-gpu=memory.used --format=csv
