#!/bin/bash
# r2l: final single-GPU state of round 1: smoke, gpu tests, default bench, launch list
mkdir -p gpurun_out
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/smoke.log
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/pytest_gpu.log
timeout 600 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"; cut -c1-400 gpurun_out/bench.json; tail -3 gpurun_out/bench.err
timeout 900 python bench.py --workload c5 --steps 4 --warmup 3 > gpurun_out/bench_c5.json 2> gpurun_out/bench_c5.err; echo "c5 rc=$?"
CMD="python bench.py --steps 1 --warmup 1 --spp 8 --no-cpu-baseline"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_launch.log 2>&1; echo "ncu launch rc=$?"
python - <<PY
import json
for f in ("bench","bench_c5"):
    for l in open(f"gpurun_out/{f}.json"):
        if l.startswith("{"):
            d=json.loads(l); print(f,"value",round(d["value"],1),"e2e",round(d["e2e"]["value"],1),"ms/step",round(d["ms_per_step"],2),"build",round(d["build"]["ms"],2),"Mprims/s",round(d["build"]["mprims_per_s"]),"frac",round(d["roofline"]["frac"],3))
PY
