set -x
mkdir -p gpurun_out
nproc
BWAGPU_BLOCKING_SYNC=${BLK:-1} timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus 8 --steps 2 --warmup 1 --no-cpu > gpurun_out/r2_8gpu_blk.json 2> gpurun_out/r2_8gpu_blk.err; echo "8gpu rc=$?"
python - <<PY
import json
d=json.load(open("gpurun_out/r2_8gpu_blk.json"))
print("N8", round(d["value"]/1e6,2), "M reads/s; e2e", round(d["e2e"]["value"]/1e6,2), "ms/step", round(d["ms_per_step"],1))
PY
