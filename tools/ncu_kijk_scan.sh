#!/bin/bash
# scan the pointwise / column kernels of one dycore step (C128) for out-of-line fp64 division paths and low DRAM use
CMD="python bench.py --workload c128 --steps 1 --warmup 1 --no-cpu-baseline --no-e2e --no-dense"
$CMD > gpurun_out/plain_c128.log 2>&1 || { echo plain failed; exit 1; }
timeout 500 ncu --section SourceCounters --section SpeedOfLight --section InstructionStats --clock-control none --import-source on --profile-from-start off \
  --kernel-name-base demangled -k "regex:k_ijk|k_ij<|k_ij_occ" -c 190 -f -o gpurun_out/r01h_kijk_scan $CMD > gpurun_out/ncu_kijk.log 2>&1
echo rc=$?
