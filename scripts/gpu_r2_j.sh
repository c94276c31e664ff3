set -x
mkdir -p gpurun_out
python -m pytest tests/test_seed_gpu.py tests/test_chain_gpu.py tests/test_memchunk_gpu.py tests/test_mem_scale_gpu.py tests/test_emf_gpu.py -m gpu -x -q > gpurun_out/r2j_tests.log 2>&1; tail -3 gpurun_out/r2j_tests.log
bash scripts/gpu_ab.sh r2j base _coop0
