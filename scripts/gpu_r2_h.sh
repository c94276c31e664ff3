set -x
mkdir -p gpurun_out
timeout 600 python bench.py --workload peaks > gpurun_out/r2h_peaks.json 2> gpurun_out/r2h_peaks.err; echo "peaks rc=$?"; cat gpurun_out/r2h_peaks.json; tail -3 gpurun_out/r2h_peaks.err
bash scripts/gpu_ab.sh r2h base _k4 _k8
