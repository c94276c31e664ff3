set -x
mkdir -p gpurun_out
python -m pytest tests/test_chain_gpu.py tests/test_memchunk_gpu.py tests/test_mem_scale_gpu.py tests/test_ert_gpu.py -m gpu -x -q > gpurun_out/r2f_tests.log 2>&1; tail -3 gpurun_out/r2f_tests.log
timeout 900 python bench.py --reads 4000000 --steps 2 --warmup 1 --no-cpu > gpurun_out/r2f_bench.json 2> gpurun_out/r2f_bench.err; echo "rc=$?"; tail -2 gpurun_out/r2f_bench.err
python - <<'PY'
import json
d = json.load(open("gpurun_out/r2f_bench.json"))
print(d["value"], d["e2e"]["value"], d["ms_per_step"])
for k in d["roofline"]["kernels"]: print(k["kernel"], round(k["ms"],1), k.get("achieved"), k.get("frac"))
PY
