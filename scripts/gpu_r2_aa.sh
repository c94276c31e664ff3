# round 2, call AA: SW cell rebalanced between the ALU and FMA pipes
set -x
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_bsw_gpu.py tests/test_c2a_gpu.py tests/test_memchunk_gpu.py -x -q -m gpu > gpurun_out/r2aa_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2aa_pytest.log
source scripts/gpu_ab2.sh
run_var r2aa base base "" "--reads 5000000 --chunk 2500000 --contexts 1"
run_var r2aa nofma _nofma "" "--reads 5000000 --chunk 2500000 --contexts 1"
run_var r2aa fma2 _fma2 "" "--reads 5000000 --chunk 2500000 --contexts 1"
BWAGPU_LIB=$PWD/bwa_mem_scale_b200/libbwagpu_fma2.so timeout 600 python -m pytest tests/test_bsw_gpu.py -x -q -m gpu > gpurun_out/r2aa_pytest_fma2.log 2>&1; echo "pytest fma2 rc=$?"; tail -2 gpurun_out/r2aa_pytest_fma2.log
