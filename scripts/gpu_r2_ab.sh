# round 2, call AB: chain build with the heavy reads placed hl per warp
set -x
mkdir -p gpurun_out
BWAGPU_LIB=$PWD/bwa_mem_scale_b200/libbwagpu_chmix.so timeout 600 python -m pytest tests/test_chain_gpu.py tests/test_memchunk_gpu.py -x -q -m gpu > gpurun_out/r2ab_pytest.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/r2ab_pytest.log
source scripts/gpu_ab2.sh
run_var r2ab base base "" "--reads 5000000 --chunk 2500000 --contexts 1"
run_var r2ab mix4 _chmix "BWAGPU_CHAIN_HEAVY_LANES=4" "--reads 5000000 --chunk 2500000 --contexts 1"
run_var r2ab mix8 _chmix "BWAGPU_CHAIN_HEAVY_LANES=8" "--reads 5000000 --chunk 2500000 --contexts 1"
run_var r2ab mix16 _chmix "BWAGPU_CHAIN_HEAVY_LANES=16" "--reads 5000000 --chunk 2500000 --contexts 1"
run_var r2ab mix2 _chmix "BWAGPU_CHAIN_HEAVY_LANES=2" "--reads 5000000 --chunk 2500000 --contexts 1"
