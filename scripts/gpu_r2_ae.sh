# round 2, call AE: fm_count through an indexable table in shared memory
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_seed_gpu.py tests/test_fmbuild_gpu.py tests/test_memchunk_gpu.py tests/test_ert_gpu.py -x -q -m gpu > gpurun_out/r2ae_pytest.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/r2ae_pytest.log
source scripts/gpu_ab2.sh
A="--reads 5000000 --chunk 2500000 --contexts 1"
run_var r2ae base base "" "$A"
run_var r2ae noct _noct "" "$A"
