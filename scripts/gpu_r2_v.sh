# round 2, call V: SW kernel v3 parity + timing; chain build in work order
set -x
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_bsw_gpu.py tests/test_c2a_gpu.py tests/test_chain_gpu.py tests/test_memchunk_gpu.py tests/test_mem_scale_gpu.py tests/test_zz_no_alloc.py tests/test_ert_gpu.py -x -q -m gpu > gpurun_out/r2v_pytest.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/r2v_pytest.log
source scripts/gpu_ab2.sh
run_var r2v base base "" "--chunk 2000000"
run_var r2v chns _chns "" "--chunk 2000000"
