# round 2, call AG: blocking host waits, one GPU (no contention expected)
set -x
mkdir -p gpurun_out
source scripts/gpu_ab2.sh
A="--reads 10000000 --chunk 2500000 --contexts 2"
run_var r2ag spin base "BWAGPU_BLOCKING_SYNC=0" "$A"
run_var r2ag block base "BWAGPU_BLOCKING_SYNC=1" "$A"
