# usage: gpu_ab.sh <tag> <variant suffixes...>  ("" = the default build): config-3 bench, 2 M reads, per-kernel ms
set -x
mkdir -p gpurun_out
TAG=$1; shift
for V in "$@"; do
  LIB=bwa_mem_scale_b200/libbwagpu$V.so
  [ "$V" = "base" ] && LIB=bwa_mem_scale_b200/libbwagpu.so
  BWAGPU_LIB=$PWD/$LIB timeout 600 python bench.py --reads 2000000 --steps 2 --warmup 1 --no-cpu > gpurun_out/${TAG}_$V.json 2> gpurun_out/${TAG}_$V.err; echo "$V rc=$?"
  python - <<PY
import json
d = json.load(open("gpurun_out/${TAG}_$V.json"))
print("VAR [$V]", round(d["value"]/1e6,3), "Mreads/s", " ".join(f'{k["kernel"].split("_kernel")[0]}={k["ms"]:.1f}' for k in d["roofline"]["kernels"]))
PY
done
