# usage: source scripts/gpu_ab2.sh; run_var <tag> <name> <lib suffix or "base"> "<ENV=.. ...>" "<extra bench args>"
# config-3 bench, 2 M reads by default, per-kernel ms
run_var() {
  local TAG=$1 NAME=$2 V=$3 ENVS=$4 EXTRA=$5
  local LIB=bwa_mem_scale_b200/libbwagpu$V.so
  [ "$V" = "base" ] && LIB=bwa_mem_scale_b200/libbwagpu.so
  env $ENVS BWAGPU_LIB=$PWD/$LIB timeout 600 python bench.py --reads 2000000 --steps 2 --warmup 1 --no-cpu $EXTRA > gpurun_out/${TAG}_$NAME.json 2> gpurun_out/${TAG}_$NAME.err; echo "$NAME rc=$?"
  python - <<PY
import json
d = json.load(open("gpurun_out/${TAG}_$NAME.json"))
print("VAR [$NAME]", round(d["value"]/1e6,3), "Mreads/s e2e", round(d["e2e"]["value"]/1e6,3), " ".join(f'{k["kernel"].split("_kernel")[0]}={k["ms"]:.1f}' for k in d["roofline"]["kernels"]))
PY
}
