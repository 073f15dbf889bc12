#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/pytest_gpu.log
run() { echo "== $*"; timeout 300 python bench.py --steps 2 --warmup 1 --spp 16 --no-cpu-baseline "$@" 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); r=d['roofline']; print('value',round(d['value'],1),'ms/step',round(d['ms_per_step'],1),'trace_ms',round(r['trace_ms_per_step'],1),'frac',round(r['frac'],3),'build',round(d['build']['ms'],2))
    elif 'rror' in l: print(l.strip())
"; }
run
run --opt wide=0
run --opt t1=8 --opt t2=8 --opt fetch_min=8
run --opt t1=20 --opt t2=20 --opt fetch_min=16
run --opt t1=0 --opt t2=0 --opt fetch_min=32
run --opt extend_blocks_per_sm=5
