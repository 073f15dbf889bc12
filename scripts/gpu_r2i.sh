#!/bin/bash
# r2i: full ncu of the shading kernels
mkdir -p gpurun_out
CMD="python bench.py --steps 1 --warmup 1 --spp 8 --no-cpu-baseline"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'k_shade' -s 1 -c 3 -o gpurun_out/shade_r2i -f $CMD > gpurun_out/ncu_full.log 2>&1; echo "ncu full rc=$?"
