#!/bin/bash
# remap column kernel compiled for 8 / 10 / 12 resident CTAs per SM; fvtp2d_march rows per chunk; C384
run() { env "$@" python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-e2e > gpurun_out/sweep_tmp.json 2>/dev/null; echo -n "$* "; python tools/benchline.py < gpurun_out/sweep_tmp.json; }
run PACE_B200_REMAP_OCC=8
run PACE_B200_REMAP_OCC=10
run PACE_B200_REMAP_OCC=12
run PACE_B200_FVTP2D_ROWS=96
run PACE_B200_FVTP2D_ROWS=192
