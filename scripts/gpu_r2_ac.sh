# round 2, call AC: SW windows read in place (no sequence arenas), finer SW tiers, 16-byte prev-list entries, L2 fetch granularity
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_seed_gpu.py tests/test_bsw_gpu.py tests/test_c2a_gpu.py tests/test_chain_gpu.py tests/test_memchunk_gpu.py tests/test_mem_scale_gpu.py tests/test_emf_gpu.py tests/test_ert_gpu.py tests/test_zz_no_alloc.py -x -q -m gpu > gpurun_out/r2ac_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2ac_pytest.log
timeout 600 python -m pytest tests/test_sam_gpu.py -x -q -m gpu -k "not ert" > gpurun_out/r2ac_pytest_sam.log 2>&1; echo "pytest sam rc=$?"; tail -2 gpurun_out/r2ac_pytest_sam.log
source scripts/gpu_ab2.sh
A="--reads 5000000 --chunk 2500000 --contexts 1"
run_var r2ac base base "" "$A"
run_var r2ac arena _arena "" "$A"
run_var r2ac tiers4 _tiers4 "" "$A"
run_var r2ac noprev16 _noprev16 "" "$A"
run_var r2ac l2f32 base "BWAGPU_L2_FETCH=32" "$A"
run_var r2ac chunk5m base "" "--reads 10000000 --chunk 5000000 --contexts 1"
