set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_sam_gpu.py -m gpu -x -q -k ert > gpurun_out/r2r_ert_sam.log 2>&1; echo "rc=$?"; tail -30 gpurun_out/r2r_ert_sam.log
ls /dev/shm | head
