# round 2, call Z: SW kernel line prefetch + row-end shortcut (parity + timing); harvested mate-rescue jobs bench
set -x
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_bsw_gpu.py tests/test_c2a_gpu.py tests/test_chain_gpu.py tests/test_memchunk_gpu.py tests/test_ksw_gpu.py -x -q -m gpu > gpurun_out/r2z_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2z_pytest.log
source scripts/gpu_ab2.sh
run_var r2z base base "" "--reads 5000000 --chunk 2500000 --contexts 1"
timeout 600 python bench.py --workload ksw --harvested 1 --pairs 2000000 > gpurun_out/r2z_ksw_harvested.json 2> gpurun_out/r2z_ksw_harvested.err; echo "ksw rc=$?"
timeout 600 python bench.py --workload ksw --harvested 1 --pairs 2000000 --impl reference > gpurun_out/r2z_ksw_harvested_ref.json 2> gpurun_out/r2z_ksw_harvested_ref.err; echo "ksw ref rc=$?"
head -c 600 gpurun_out/r2z_ksw_harvested.json; echo; head -c 400 gpurun_out/r2z_ksw_harvested_ref.json
