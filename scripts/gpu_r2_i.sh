set -x
mkdir -p gpurun_out
timeout 600 python bench.py --workload peaks > gpurun_out/r2i_peaks.json 2> gpurun_out/r2i_peaks.err; echo "peaks rc=$?"; python -c "
import json; d=json.load(open('gpurun_out/r2i_peaks.json'))
for k,v in d['bext_probe'].items(): print(k, v)"; tail -3 gpurun_out/r2i_peaks.err
