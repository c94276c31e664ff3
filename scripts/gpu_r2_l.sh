set -x
mkdir -p gpurun_out
python -m pytest tests/test_emf_gpu.py -m gpu -x -q > gpurun_out/r2l_emf.log 2>&1; tail -30 gpurun_out/r2l_emf.log
