# round 2, call B: worker-level entry + SAM identity through the compiled shim, then the whole suite
set -x
mkdir -p gpurun_out
python -m pytest tests/test_memchunk_gpu.py tests/test_sam_gpu.py -m gpu -x -q > gpurun_out/r2b_new.log 2>&1; tail -30 gpurun_out/r2b_new.log
python -m pytest tests -m gpu -x -q > gpurun_out/r2b_pytest_gpu.log 2>&1; tail -3 gpurun_out/r2b_pytest_gpu.log
