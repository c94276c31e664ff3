# round 2, call U: paired CP_OCC layout (half-record backward extension) + two backward-row items per step
set -x
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_seed_gpu.py tests/test_fmbuild_gpu.py tests/test_memchunk_gpu.py tests/test_emf_gpu.py tests/test_ert_gpu.py tests/test_mem_scale_gpu.py -x -q -m gpu > gpurun_out/r2u_pytest.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/r2u_pytest.log
timeout 600 python -m pytest tests/test_sam_gpu.py -x -q -m gpu -k "not ert" > gpurun_out/r2u_pytest_sam.log 2>&1; echo "pytest sam rc=$?"; tail -3 gpurun_out/r2u_pytest_sam.log
source scripts/gpu_ab2.sh
run_var r2u base base "" "--chunk 2000000"
run_var r2u mb5 _mb5 "" "--chunk 2000000"
run_var r2u nopair2 _nopair2 "" "--chunk 2000000"
run_var r2u nopair2_sb8 _nopair2_sb8 "" "--chunk 2000000"
run_var r2u mb5_sb8 _mb5_sb8 "" "--chunk 2000000"
