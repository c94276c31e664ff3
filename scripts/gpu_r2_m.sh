set -x
mkdir -p gpurun_out
free -g | head -2
timeout 1500 python bench.py --steps 3 --warmup 2 > gpurun_out/r2m_full.json 2> gpurun_out/r2m_full.err; echo "full rc=$?"; grep "bench\]" gpurun_out/r2m_full.err; tail -3 gpurun_out/r2m_full.err
timeout 1500 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r2m_ref.json 2> gpurun_out/r2m_ref.err; echo "ref rc=$?"; tail -2 gpurun_out/r2m_ref.err
python - <<'PY'
import json
for f in ("gpurun_out/r2m_full.json", "gpurun_out/r2m_ref.json"):
    d = json.load(open(f))
    print(f, d["value"], d["e2e"]["value"], d["ms_per_step"], d.get("cpu_baseline"), d.get("setup"))
    if "roofline" in d:
        for k in d["roofline"]["kernels"]: print("  ", k["kernel"], round(k["ms"],1), k.get("achieved"), k.get("frac"))
        print(d["work"], d["gcups"])
PY
