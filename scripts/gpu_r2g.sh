#!/bin/bash
# r2g: primary-visibility split (rt_render_primary / rt_scene_set_primary_hits) on two GPUs
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log
for sp in 1 0; do
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 2954$sp bench.py --gpus 2 --steps 3 --warmup 3 --split-primary $sp > gpurun_out/bench_n2_sp$sp.json 2> gpurun_out/bench_n2_sp$sp.err; echo "n2 split=$sp rc=$?"; grep -v "OMP_NUM\|\*\*\*" gpurun_out/bench_n2_sp$sp.err | tail -2
done
python - <<PY
import json
for f in ("sp1","sp0"):
    for l in open(f"gpurun_out/bench_n2_{f}.json"):
        if l.startswith("{"):
            d=json.loads(l); print(f,"value",round(d["value"],1),"e2e",round(d["e2e"]["value"],1),"ms/step",round(d["ms_per_step"],2),"segs/step",d["segments_per_step"])
PY
