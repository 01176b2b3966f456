#!/bin/bash
# round-end evidence: launch list of one step and DRAM traffic of every fvtp2d_march launch (C384)
CMD="python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-e2e"
$CMD > gpurun_out/plain_c384.log 2>&1 || { echo plain failed; exit 1; }
ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r01h_launches_c384.csv $CMD > gpurun_out/ncu_launches.log 2>&1
echo launches rc=$?
ncu --clock-control none --profile-from-start off --kernel-name-base demangled -k regex:fvtp2d_march -c 60 --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --csv --log-file gpurun_out/r01h_march_traffic.csv $CMD > gpurun_out/ncu_traffic.log 2>&1
echo traffic rc=$?
