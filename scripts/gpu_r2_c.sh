set -x
mkdir -p gpurun_out
BWAGPU_VERBOSE=1 python -m pytest tests/test_fmbuild_gpu.py -m gpu -x -q -s > gpurun_out/r2c_fmbuild.log 2>&1; tail -40 gpurun_out/r2c_fmbuild.log
