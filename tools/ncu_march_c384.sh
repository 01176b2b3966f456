#!/bin/bash
# fvtp2d_march at the bench workload (C384): scheduler / warp-state / source counters of one launch of each order
set -u
CMD="python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-e2e"
$CMD > gpurun_out/plain_c384.log 2>&1 || { echo "plain run failed"; exit 1; }
cap() {
  timeout 400 ncu --section WarpStateStats --section SchedulerStats --section SpeedOfLight --section Occupancy --section SourceCounters --section InstructionStats \
      --clock-control none --import-source on --profile-from-start off --kernel-name-base demangled \
      -k "regex:$2" -s "$3" -c 1 -f -o "gpurun_out/r01g_$1" $CMD > "gpurun_out/ncu_$1.log" 2>&1
  echo "$1 rc=$?"
}
cap march66_c384 'k_fvtp2d_march<.int.6' 0
cap march88_c384 'k_fvtp2d_march<.int.8' 0
