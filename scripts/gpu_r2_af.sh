# round 2, call AF: backward rows of <= 2 entries go on from registers (no prev-list load)
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_seed_gpu.py tests/test_memchunk_gpu.py tests/test_mem_scale_gpu.py tests/test_emf_gpu.py -x -q -m gpu > gpurun_out/r2af_pytest.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/r2af_pytest.log
source scripts/gpu_ab2.sh
A="--reads 5000000 --chunk 2500000 --contexts 1"
run_var r2af base base "" "$A"
run_var r2af nofwd _nofwd "" "$A"
