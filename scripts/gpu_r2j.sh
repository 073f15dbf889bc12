#!/bin/bash
# r2j: k_shade at 3 blocks/SM (40 registers)
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/pytest_gpu.log
run() { echo "== $*"; timeout 300 python bench.py --steps 3 --warmup 2 --no-cpu-baseline "$@" 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); r=d['roofline']; print('value',round(d['value'],1),'e2e',round(d['e2e']['value'],1),'ms/step',round(d['ms_per_step'],2),'trace_ms',round(r['trace_ms_per_step'],1),'build',round(d['build']['ms'],2))
    elif 'rror' in l: print(l.strip())
"; }
run
run
