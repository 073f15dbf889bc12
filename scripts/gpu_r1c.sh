#!/bin/bash
mkdir -p gpurun_out
run() { echo "== $*"; timeout 300 python bench.py --steps 2 --warmup 1 --spp 16 --no-cpu-baseline "$@" 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); r=d['roofline']; print('value',round(d['value'],1),'ms/step',round(d['ms_per_step'],1),'trace_ms',round(r['trace_ms_per_step'],1),'frac',round(r['frac'],3),'build',round(d['build']['ms'],2))
    elif 'rror' in l: print(l.strip())
"; }
run --opt cull=0
run --opt cull=1
run --opt cull=0 --leaf-max 1
run --opt cull=1 --leaf-max 1
CMD="python bench.py --steps 1 --warmup 1 --spp 8 --no-cpu-baseline"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_extend_wide -s 3 -c 2 -o gpurun_out/k_extend_wide_cull -f $CMD > gpurun_out/ncu_full.log 2>&1; echo "ncu rc=$?"
