# round 2, call S: seed_collect_kernel — dynamic read fetch vs static stride, occupancy variants, chunk size
set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_seed_gpu.py tests/test_memchunk_gpu.py -x -q -m gpu > gpurun_out/r2s_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2s_pytest.log
bash scripts/gpu_ab.sh r2s base _static _mb8 _mb6
BWAGPU_LIB=$PWD/bwa_mem_scale_b200/libbwagpu.so timeout 600 python bench.py --reads 5000000 --chunk 2500000 --steps 2 --warmup 1 --no-cpu > gpurun_out/r2s_chunk25.json 2> gpurun_out/r2s_chunk25.err; echo "chunk rc=$?"
python - <<PY
import json
d = json.load(open("gpurun_out/r2s_chunk25.json"))
print("VAR [chunk 2.5M, 5M reads]", round(d["value"]/1e6,3), "Mreads/s", " ".join(f'{k["kernel"].split("_kernel")[0]}={k["ms"]:.1f}' for k in d["roofline"]["kernels"]))
PY
