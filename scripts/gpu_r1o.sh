#!/bin/bash
# r1o: two GPUs: NCCL/NVLink versions of the multi-GPU tests, bench at N=2 with both exchange variants
mkdir -p gpurun_out
nvidia-smi -L; nvidia-smi topo -m 2>/dev/null | head -6
timeout 600 python -m pytest tests/test_gpu_multigpu.py -m gpu -x -q > gpurun_out/pytest_gpu2.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_gpu2.log
for ex in peer nccl; do
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 3 --warmup 3 --exchange $ex > gpurun_out/bench_n2_$ex.json 2> gpurun_out/bench_n2_$ex.err; echo "bench n2 $ex rc=$?"; tail -c 1800 gpurun_out/bench_n2_$ex.json; tail -3 gpurun_out/bench_n2_$ex.err
done
