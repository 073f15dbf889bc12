#!/bin/bash
# r1y: batch size sweep
mkdir -p gpurun_out
run() { echo "== $*"; timeout 300 python bench.py --steps 2 --warmup 2 --no-cpu-baseline "$@" 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); r=d['roofline']; print('value',round(d['value'],1),'e2e',round(d['e2e']['value'],1),'ms/step',round(d['ms_per_step'],1),'trace_ms',round(r['trace_ms_per_step'],1),'frac',round(r['frac'],3),'build',round(d['build']['ms'],2))
    elif 'rror' in l: print(l.strip())
"; }
run --batch-spp 2
run --batch-spp 8
run --batch-spp 16
