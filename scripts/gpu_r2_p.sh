set -x
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/r2p_pytest_gpu.log 2>&1; tail -3 gpurun_out/r2p_pytest_gpu.log
bash scripts/gpu_ab.sh r2p base
