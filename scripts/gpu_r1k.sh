#!/bin/bash
# r1k: threshold sweep after octant binning
mkdir -p gpurun_out
run() { echo "== $*"; timeout 300 python bench.py --steps 2 --warmup 2 --no-cpu-baseline "$@" 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); r=d['roofline']; print('value',round(d['value'],1),'ms/step',round(d['ms_per_step'],1),'trace_ms',round(r['trace_ms_per_step'],1),'frac',round(r['frac'],3),'build',round(d['build']['ms'],2))
    elif 'rror' in l: print(l.strip())
"; }
run --opt t1=16 --opt t2=4
run --opt t1=20 --opt t2=8
run --opt t1=20 --opt t2=4
run --opt t1=24 --opt t2=8
run --opt t1=16 --opt t2=12
run --opt t1=16 --opt t2=8 --opt fetch_min=2
run --opt t1=16 --opt t2=8 --opt fetch_min=1
run --opt t1=16 --opt t2=8 --opt extend_blocks_per_sm=7
