# round 2, call W: worker contexts over one index; cooperative read staging
set -x
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_memchunk_gpu.py tests/test_seed_gpu.py tests/test_mem_scale_gpu.py tests/test_chain_gpu.py -x -q -m gpu > gpurun_out/r2w_pytest.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/r2w_pytest.log
source scripts/gpu_ab2.sh
run_var r2w base base "" "--reads 5000000 --chunk 2500000"
run_var r2w nostc _nostc "" "--reads 5000000 --chunk 2500000"
run_var r2w ctx2 base "" "--reads 5000000 --chunk 1250000 --contexts 2"
run_var r2w ctx2_c25 base "" "--reads 10000000 --chunk 2500000 --contexts 2"
run_var r2w ctx3 base "" "--reads 7500000 --chunk 1250000 --contexts 3"
