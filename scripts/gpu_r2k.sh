#!/bin/bash
# r2k: occupancy variants of k_sort_pass / k_collapse (launch list only)
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_build.py -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/pytest_gpu.log
CMD="python bench.py --steps 1 --warmup 1 --spp 4 --no-cpu-baseline"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_launch.log 2>&1; echo "ncu launch rc=$?"
CMD="python bench.py --workload c5 --steps 1 --warmup 1"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 150 --csv --log-file gpurun_out/launches_c5.csv $CMD > gpurun_out/ncu_launch_c5.log 2>&1; echo "ncu launch c5 rc=$?"
