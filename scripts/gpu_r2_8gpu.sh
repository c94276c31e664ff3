set -x
mkdir -p gpurun_out
nvidia-smi -L | head -8; free -g | head -2; nproc
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus 8 --steps 2 --warmup 1 --no-cpu > gpurun_out/r2_8gpu.json 2> gpurun_out/r2_8gpu.err; echo "8gpu rc=$?"
tail -c 1500 gpurun_out/r2_8gpu.json | head -c 700; echo; tail -5 gpurun_out/r2_8gpu.err
