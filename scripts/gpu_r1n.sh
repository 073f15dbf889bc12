#!/bin/bash
# r1n: peer exchange test on one GPU (two processes, CUDA IPC on the shared device) + bench
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/pytest_gpu.log
timeout 300 python bench.py --steps 2 --warmup 2 --no-cpu-baseline 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); r=d['roofline']; print('value',round(d['value'],1),'ms/step',round(d['ms_per_step'],1),'trace_ms',round(r['trace_ms_per_step'],1),'frac',round(r['frac'],3),'build',round(d['build']['ms'],2))
    elif 'rror' in l: print(l.strip())
"
