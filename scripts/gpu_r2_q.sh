set -x
mkdir -p gpurun_out
python -m pytest tests/test_seed_gpu.py -m gpu -x -q 2>&1 | tail -2
BWAGPU_VERBOSE=1 bash scripts/gpu_ab.sh r2q base
grep "outgrow" gpurun_out/r2q_base.err | head -5
