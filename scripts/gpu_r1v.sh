#!/bin/bash
# r1v: k_flatten (block-level instance lookup, 4 prims/thread), vectorised k_spheres_advect
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log
timeout 300 python bench.py --steps 2 --warmup 2 --no-cpu-baseline 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); r=d['roofline']; print('c4 value',round(d['value'],1),'ms/step',round(d['ms_per_step'],1),'build',round(d['build']['ms'],2))
"
CMD="python bench.py --workload c5 --steps 1 --warmup 1"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 150 --csv --log-file gpurun_out/launches_c5.csv $CMD > gpurun_out/ncu_launch_c5.log 2>&1; echo "ncu launch c5 rc=$?"
timeout 900 python bench.py --workload c5 --steps 3 --warmup 3 > gpurun_out/bench_c5.json 2>gpurun_out/bench_c5.err; python - <<PY
import json
for l in open("gpurun_out/bench_c5.json"):
    if l.startswith("{"):
        d=json.loads(l); print("c5 value",round(d["value"],1),"e2e",round(d["e2e"]["value"],1),"ms/step",round(d["ms_per_step"],2),"build",round(d["build"]["ms"],2),"Mprims/s",round(d["build"]["mprims_per_s"]),"fps",round(d["frames_per_s"],1))
PY
