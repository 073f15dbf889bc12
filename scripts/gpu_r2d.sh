#!/bin/bash
# r2d: collapse micro-optimisations (cached areas, branch-free slot fallback)
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log
for w in c4 c5; do
timeout 900 python bench.py --workload $w --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_$w.json 2>gpurun_out/bench_$w.err; echo "$w rc=$?"
done
python - <<PY
import json
for f in ("c4","c5"):
    for l in open(f"gpurun_out/bench_{f}.json"):
        if l.startswith("{"):
            d=json.loads(l); print(f,"value",round(d["value"],1),"e2e",round(d["e2e"]["value"],1),"ms/step",round(d["ms_per_step"],2),"build",round(d["build"]["ms"],2),"Mprims/s",round(d["build"]["mprims_per_s"]))
PY
