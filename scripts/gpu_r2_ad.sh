# round 2, call AD: chunk size x contexts with the in-place SW windows
set -x
mkdir -p gpurun_out
source scripts/gpu_ab2.sh
run_var r2ad k1_c25 base "" "--reads 10000000 --chunk 2500000 --contexts 1"
run_var r2ad k2_c25 base "" "--reads 10000000 --chunk 2500000 --contexts 2"
run_var r2ad k1_c5 base "" "--reads 10000000 --chunk 5000000 --contexts 1"
run_var r2ad k2_c5 base "" "--reads 10000000 --chunk 5000000 --contexts 2"
run_var r2ad k2_c33 base "" "--reads 10000000 --chunk 3400000 --contexts 2"
