#!/bin/bash
# r2m / r2p: C4 only, final code: scaling on one 8-GPU box: C4 (spp split, peer exchange) at N=1,2,4,8 and C5 (frame-parallel) at N=1,8
mkdir -p gpurun_out
nvidia-smi -L | head -8; free -g | head -2
for n in 1 2 4 8; do
  if [ $n -eq 1 ]; then L="python"; else L="python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2951$n"; fi
  timeout 600 $L bench.py --gpus $n --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/scale_c4_n$n.json 2> gpurun_out/scale_c4_n$n.err; echo "c4 n=$n rc=$?"
  python - <<PY
import json
for l in open("gpurun_out/scale_c4_n$n.json"):
    if l.startswith("{"):
        d=json.loads(l); print("  value",round(d["value"],1),"e2e",round(d["e2e"]["value"],1),"ms/step",round(d["ms_per_step"],2),"build",round(d["build"]["ms"],2),"frac",round(d["roofline"]["frac"],3))
PY
done
