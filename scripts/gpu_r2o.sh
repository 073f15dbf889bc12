#!/bin/bash
# r2o: RNG sample stride 32 (was 64): full gpu suite + bench
mkdir -p gpurun_out
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/smoke.log
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/pytest_gpu.log
timeout 600 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"; tail -3 gpurun_out/bench.err
python - <<PY
import json
for f in ("bench",):
    for l in open(f"gpurun_out/{f}.json"):
        if l.startswith("{"):
            d=json.loads(l); print(f,"value",round(d["value"],1),"e2e",round(d["e2e"]["value"],1),"ms/step",round(d["ms_per_step"],2),"build",round(d["build"]["ms"],2),"Mprims/s",round(d["build"]["mprims_per_s"]),"frac",round(d["roofline"]["frac"],3),"cpu",d["cpu_baseline"]["value"])
PY
