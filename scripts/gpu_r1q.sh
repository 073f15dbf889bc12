#!/bin/bash
# r1q: collapse without divisions, two-kernel refit (option refit_mode=1)
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_gpu.log
run() { echo "== $*"; timeout 300 python bench.py --steps 2 --warmup 2 --no-cpu-baseline "$@" 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); r=d['roofline']; print('value',round(d['value'],1),'ms/step',round(d['ms_per_step'],1),'trace_ms',round(r['trace_ms_per_step'],1),'frac',round(r['frac'],3),'build',round(d['build']['ms'],2))
    elif 'rror' in l: print(l.strip())
"; }
run
run --opt refit_mode=1
CMD="python bench.py --steps 1 --warmup 1 --spp 4 --no-cpu-baseline --opt refit_mode=1"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_launch.log 2>&1; echo "ncu launch rc=$?"
