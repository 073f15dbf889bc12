#!/bin/bash
# r2a: two GPUs: tests, C4 at N=2 (auto batch, RNG repositioning on the side stream, pipelined e2e inputs)
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log
timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; echo "n1 rc=$?"
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29531 bench.py --gpus 2 --steps 3 --warmup 3 > gpurun_out/bench_n2.json 2> gpurun_out/bench_n2.err; echo "n2 rc=$?"; tail -2 gpurun_out/bench_n2.err
python - <<PY
import json
for f in ("n1","n2"):
    for l in open(f"gpurun_out/bench_{f}.json"):
        if l.startswith("{"):
            d=json.loads(l); print(f,"value",round(d["value"],1),"e2e",round(d["e2e"]["value"],1),"ms/step",round(d["ms_per_step"],2),"e2e ms",round(d["e2e"]["ms_per_step"],2),"build",round(d["build"]["ms"],2),"frac",round(d["roofline"]["frac"],3))
PY
