set -x
mkdir -p gpurun_out
nvidia-smi -L; nproc; free -g | head -2
timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus 2 --steps 3 --warmup 2 --no-cpu > gpurun_out/r2n_2gpu.json 2> gpurun_out/r2n_2gpu.err; echo "2gpu rc=$?"; tail -4 gpurun_out/r2n_2gpu.err
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29518 bench.py --impl reference --gpus 2 --steps 1 --warmup 1 --cpu-reads-per-thread 20000 > gpurun_out/r2n_2gpu_ref.json 2> gpurun_out/r2n_2gpu_ref.err; echo "ref rc=$?"; tail -2 gpurun_out/r2n_2gpu_ref.err
python - <<'PY'
import json
for f in ("gpurun_out/r2n_2gpu.json", "gpurun_out/r2n_2gpu_ref.json"):
    d = json.load(open(f)); print(f, d["n_gpus"], d["value"], d["e2e"]["value"], d["ms_per_step"], d.get("setup"))
PY
