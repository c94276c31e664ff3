# round 2, call T: BSW_V2 parity + A/B; seed kernel knobs on top of the dynamic read fetch
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_bsw_gpu.py tests/test_c2a_gpu.py tests/test_memchunk_gpu.py tests/test_mem_scale_gpu.py tests/test_zz_no_alloc.py tests/test_ksw_gpu.py -x -q -m gpu > gpurun_out/r2t_pytest.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/r2t_pytest.log
source scripts/gpu_ab2.sh
run_var r2t base base "" ""
run_var r2t bswv1 _bswv1 "" ""
run_var r2t sb8 _sb8 "" ""
run_var r2t sb2 _sb2 "" ""
run_var r2t ld0 _ld0 "" ""
run_var r2t ld2 _ld2 "" ""
run_var r2t pk4 _pk4 "" ""
run_var r2t l2f32 base "BWAGPU_L2_FETCH=32" ""
run_var r2t l2f128 base "BWAGPU_L2_FETCH=128" ""
run_var r2t bps4 base "BWAGPU_SEED_BLOCKS_PER_SM=4" ""
run_var r2t bps3 base "BWAGPU_SEED_BLOCKS_PER_SM=3" ""
run_var r2t chunk2m base "" "--chunk 2000000"
run_var r2t r10m_c25 base "" "--reads 10000000 --chunk 2500000"
run_var r2t r10m_c5 base "" "--reads 10000000 --chunk 5000000"
