# round 2, call X: worker contexts over one index (after the shared-memory-bound fix)
set -x
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_memchunk_gpu.py tests/test_ksw_gpu.py tests/test_seed_gpu.py -x -q -m gpu > gpurun_out/r2x_pytest.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/r2x_pytest.log
source scripts/gpu_ab2.sh
run_var r2x k1 base "" "--reads 10000000 --chunk 2500000"
run_var r2x k2 base "" "--reads 10000000 --chunk 2500000 --contexts 2"
run_var r2x k2_c125 base "" "--reads 10000000 --chunk 1250000 --contexts 2"
run_var r2x k3_c125 base "" "--reads 10000000 --chunk 1250000 --contexts 3"
