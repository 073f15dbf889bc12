#!/bin/bash
# r1r: full ncu of the traversal kernel after octant binning (bounce launches)
mkdir -p gpurun_out
CMD="python bench.py --steps 1 --warmup 1 --spp 8 --no-cpu-baseline"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_extend_wide -s 3 -c 2 -o gpurun_out/k_extend_wide_r1r -f $CMD > gpurun_out/ncu_full.log 2>&1; echo "ncu full rc=$?"
