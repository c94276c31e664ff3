# round 2, final evidence: the whole gpu suite, smoke, the default bench (both arms), ncu captures of the two top kernels
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/r2f2_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2f2_pytest_gpu.log
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2f2_smoke.log 2>&1; echo "smoke rc=$?"; tail -4 gpurun_out/r2f2_smoke.log
timeout 900 python bench.py --impl reference > gpurun_out/r2f2_bench_reference.json 2> gpurun_out/r2f2_bench_reference.err; echo "ref rc=$?"
timeout 900 python bench.py > gpurun_out/r2f2_bench.json 2> gpurun_out/r2f2_bench.err; echo "bench rc=$?"
python - <<PY
import json
a=json.load(open("gpurun_out/r2f2_bench.json")); b=json.load(open("gpurun_out/r2f2_bench_reference.json"))
print("BENCH", round(a["value"]/1e6,3), "e2e", round(a["e2e"]["value"]/1e6,3), "ref", round(b["value"]/1e3,1), "k reads/s; same_config", a["config"]==b["config"], "ratio e2e", round(a["e2e"]["value"]/b["value"],1))
print(" ".join(f'{k["kernel"].split("_kernel")[0]}={k["ms"]:.1f}' for k in a["roofline"]["kernels"]), "gcups", round(a["gcups"],1), "cpu", a.get("cpu_baseline",{}).get("value"))
PY
CMD="python bench.py --reads 2500000 --chunk 2500000 --contexts 1 --steps 1 --warmup 1 --no-cpu"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:seed_collect_kernel -s 0 -c 1 -f -o gpurun_out/r2f2_seed_collect_kernel $CMD > gpurun_out/r2f2_ncu_seed.log 2>&1; echo "seed ncu rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:bsw_fast_kernel -s 0 -c 1 -f -o gpurun_out/r2f2_bsw_fast_kernel $CMD > gpurun_out/r2f2_ncu_bsw.log 2>&1; echo "bsw ncu rc=$?"
