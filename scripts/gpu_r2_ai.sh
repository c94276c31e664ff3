# round 2, call AI: early-exit key search in the chain B-tree nodes
set -x
mkdir -p gpurun_out
BWAGPU_LIB=$PWD/bwa_mem_scale_b200/libbwagpu_lin.so timeout 600 python -m pytest tests/test_chain_gpu.py tests/test_memchunk_gpu.py tests/test_mem_scale_gpu.py -x -q -m gpu > gpurun_out/r2ai_pytest.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/r2ai_pytest.log
source scripts/gpu_ab2.sh
A="--reads 5000000 --chunk 2500000 --contexts 1"
run_var r2ai base base "" "$A"
run_var r2ai lin _lin "" "$A"
