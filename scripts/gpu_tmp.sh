mkdir -p gpurun_out
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29561 bench.py --gpus 2 --steps 2 --warmup 3 > gpurun_out/bench_n2_final.json 2> gpurun_out/bench_n2_final.err; echo "n2 rc=$?"; grep -v "OMP_NUM\|\*\*\*" gpurun_out/bench_n2_final.err | tail -3
python - <<PY
import json
for l in open("gpurun_out/bench_n2_final.json"):
    if l.startswith("{"):
        d=json.loads(l); print("value",round(d["value"],1),"e2e",round(d["e2e"]["value"],1),"|",d["config"]["primary_rays"],"|",d["config"]["parallelism"])
PY
