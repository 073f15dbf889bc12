#!/bin/bash
# r1z: auto batch (16 spp in flight): tests + threshold re-sweep
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log
run() { echo "== $*"; timeout 300 python bench.py --steps 2 --warmup 2 --no-cpu-baseline "$@" 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); r=d['roofline']; print('value',round(d['value'],1),'e2e',round(d['e2e']['value'],1),'ms/step',round(d['ms_per_step'],1),'trace_ms',round(r['trace_ms_per_step'],1),'frac',round(r['frac'],3),'build',round(d['build']['ms'],2))
    elif 'rror' in l: print(l.strip())
"; }
run
run --opt extend_blocks_per_sm=8
run --opt t1=12 --opt t2=4
run --opt t1=20 --opt t2=4
run --opt t1=16 --opt t2=2
run --opt fetch_min=8
run --batch-spp 32
