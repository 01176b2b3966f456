#!/bin/bash
# second pass: the kernels whose demangled names carry template arguments, and all column kernels of remap_main
set -u
CMD="python bench.py --workload c128 --steps 1 --warmup 1 --no-cpu-baseline --no-e2e"
$CMD > gpurun_out/plain_c128.log 2>&1 || { echo "plain run failed"; exit 1; }
cap() {  # name regex skip count
  ncu --set full --clock-control none --import-source on --profile-from-start off --kernel-name-base demangled \
      -k "regex:$2" -s "$3" -c "$4" -f -o "gpurun_out/r01f_$1" $CMD > "gpurun_out/ncu_$1.log" 2>&1
  echo "$1 rc=$?"
}
cap march66 'k_fvtp2d_march<.int.6' 2 1
cap march88 'k_fvtp2d_march<.int.8' 2 1
cap riem0 'k_riem<.bool.0' 1 1
cap riem1 'k_riem<.bool.1' 1 1
cap remap 'k_ij<pace::remap_main' 0 5
