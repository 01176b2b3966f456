#!/bin/bash
# remap column kernel: resident CTAs per SM (unused dynamic shared memory) vs time, C384
for kb in 0 28 37 56 75 113; do
  PACE_B200_REMAP_SMEM_KB=$kb python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-e2e > gpurun_out/sweep_$kb.json 2>/dev/null
  echo -n "smem_kb=$kb "; python tools/benchline.py < gpurun_out/sweep_$kb.json
done
CMD="python bench.py --workload c128 --steps 1 --warmup 1 --no-cpu-baseline --no-e2e"
$CMD > gpurun_out/plain_c128.log 2>&1 && ncu --set full --clock-control none --import-source on --profile-from-start off --kernel-name-base demangled -k "regex:k_ij<pace::remap_main" -s 1 -c 1 -f -o gpurun_out/r01g_remap $CMD > gpurun_out/ncu_remap2.log 2>&1
echo ncu rc=$?
