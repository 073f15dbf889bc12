#!/bin/bash
# r1x: e2e with pipelined input staging (C4 and C5, one GPU)
mkdir -p gpurun_out
timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_c4.json 2> gpurun_out/bench_c4.err; echo "c4 rc=$?"; tail -2 gpurun_out/bench_c4.err
timeout 900 python bench.py --workload c5 --steps 3 --warmup 3 > gpurun_out/bench_c5.json 2>gpurun_out/bench_c5.err; echo "c5 rc=$?"; tail -2 gpurun_out/bench_c5.err
python - <<PY
import json
for f in ("c4","c5"):
    for l in open(f"gpurun_out/bench_{f}.json"):
        if l.startswith("{"):
            d=json.loads(l); print(f,"value",round(d["value"],1),"e2e",round(d["e2e"]["value"],1),"ms/step",round(d["ms_per_step"],2),"e2e ms",round(d["e2e"]["ms_per_step"],2),"build",round(d["build"]["ms"],2))
PY
