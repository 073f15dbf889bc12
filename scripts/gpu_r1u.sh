#!/bin/bash
# r1u: ray-bin key / shade block experiments; launch list of the C5 workload
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/pytest_gpu.log
run() { echo "== $*"; timeout 300 python bench.py --steps 2 --warmup 2 --no-cpu-baseline "$@" 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); r=d['roofline']; print('value',round(d['value'],1),'ms/step',round(d['ms_per_step'],1),'trace_ms',round(r['trace_ms_per_step'],1),'frac',round(r['frac'],3),'build',round(d['build']['ms'],2))
    elif 'rror' in l: print(l.strip())
"; }
run
run --opt ray_bins=24
run --opt shade_threads=1024
run --opt ray_bins=24 --opt shade_threads=1024
CMD="python bench.py --workload c5 --steps 1 --warmup 1"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 150 --csv --log-file gpurun_out/launches_c5.csv $CMD > gpurun_out/ncu_launch_c5.log 2>&1; echo "ncu launch c5 rc=$?"
