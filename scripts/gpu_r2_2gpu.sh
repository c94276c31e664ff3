set -x
mkdir -p gpurun_out
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29543 bench.py --gpus 2 --steps 3 --warmup 2 --no-cpu > gpurun_out/r2_2gpu_weak.json 2> gpurun_out/r2_2gpu_weak.err; echo "weak rc=$?"
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29544 bench.py --gpus 2 --steps 3 --warmup 2 --no-cpu --scaling strong > gpurun_out/r2_2gpu_strong.json 2> gpurun_out/r2_2gpu_strong.err; echo "strong rc=$?"
python - <<PY
import json
for n in ("weak","strong"):
    d=json.load(open(f"gpurun_out/r2_2gpu_{n}.json")); print("N2", n, round(d["value"]/1e6,2), "M reads/s; e2e", round(d["e2e"]["value"]/1e6,2), "ms/step", round(d["ms_per_step"],1), d["scaling"], d["config"]["reads_per_gpu"])
PY
