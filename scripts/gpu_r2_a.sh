# round 2, call A: no-alloc entry point (alone, then after the whole suite), compute-sanitizer over the golden ERT tests
set -x
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,memory.total --format=csv; nproc; free -g | head -2
python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > gpurun_out/r2a_noalloc_alone.log 2>&1; tail -3 gpurun_out/r2a_noalloc_alone.log
python -m pytest tests -m gpu -x -q > gpurun_out/r2a_pytest_gpu.log 2>&1; tail -3 gpurun_out/r2a_pytest_gpu.log
timeout 900 compute-sanitizer --tool memcheck --print-limit 20 python -m pytest tests/test_ert_gpu.py -m gpu -x -q -k "not scale" > gpurun_out/r2a_ert_memcheck.log 2>&1; tail -5 gpurun_out/r2a_ert_memcheck.log
timeout 900 compute-sanitizer --tool initcheck --print-limit 20 python -m pytest tests/test_ert_gpu.py -m gpu -x -q -k "not scale" > gpurun_out/r2a_ert_initcheck.log 2>&1; tail -5 gpurun_out/r2a_ert_initcheck.log
