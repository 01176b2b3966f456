#!/bin/bash
# ncu --set full captures of the heaviest kernels of one dycore step (C128 L79, 6 subdomains), one launch each.
# Run under gpurun after the plain command has exited 0; reports land in gpurun_out/.
set -u
CMD="python bench.py --workload c128 --steps 1 --warmup 1 --no-cpu-baseline --no-e2e"
$CMD > gpurun_out/plain_c128.log 2>&1 || { echo "plain run failed"; exit 1; }
cap() {  # name regex skip
  ncu --set full --clock-control none --import-source on --profile-from-start off --kernel-name-base demangled \
      -k "regex:$2" -s "$3" -c 1 -f -o "gpurun_out/r01f_$1" $CMD > "gpurun_out/ncu_$1.log" 2>&1
  echo "$1 rc=$?"
}
cap march66 'k_fvtp2d_march<.int.6' 2
cap march88 'k_fvtp2d_march<.int.8' 2
cap a2b 'k_a2b_fused' 1
cap delnflux 'k_delnflux_fused' 1
cap riem0 'k_riem<.bool.0' 1
cap riem1 'k_riem<.bool.1' 1
cap remap 'remap_main' 0
cap d_sw_ke 'k_ijk.*d_sw' 7
