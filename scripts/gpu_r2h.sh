#!/bin/bash
# r2h: one GPU: full gpu suite (incl. the split-primary frame of the 2-process peer test over CUDA IPC)
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log
timeout 300 python bench.py --workload c5 --n-spheres 5000000 --steps 2 --warmup 1 2>&1 | tail -c 600
