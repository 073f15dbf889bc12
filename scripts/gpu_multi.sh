#!/bin/bash
# scripts/gpu_multi.sh N TAG [variant args ...] — N-GPU C4 bench lines (torchrun, one rank per GPU), one per quoted variant
# argument string ("" = defaults); for `gpurun --gpus N -- 'bash scripts/gpu_multi.sh N tag "" "--split-primary 1"'`.
set -u
N=$1; TAG=$2; shift 2
O=gpurun_out; mkdir -p $O
[ $# -eq 0 ] && set -- ""
i=0
for v in "$@"; do
  i=$((i+1))
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29500+i)) \
      bench.py --gpus $N --steps 5 --warmup 3 $v > $O/${TAG}_n${N}_v$i.json 2> $O/${TAG}_n${N}_v$i.err
  echo "== N=$N [$i] $v rc=$?"
  python - $O/${TAG}_n${N}_v$i.json <<'PY'
import json, sys
for l in open(sys.argv[1]):
    if l.startswith('{'):
        d = json.loads(l)
        print('value %.1f e2e %.1f ms/step %.2f e2e_ms %.2f h2d %s parity %s' % (d['value'], d['e2e']['value'], d['ms_per_step'], d['e2e']['ms_per_step'],
              d['e2e']['h2d_bytes_per_step'], d.get('parity_check')))
PY
  tail -2 $O/${TAG}_n${N}_v$i.err
done
