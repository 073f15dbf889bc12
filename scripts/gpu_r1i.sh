#!/bin/bash
# r1i: state verification after container re-creation: gpu tests, default bench (both arms), launch list, full ncu of the top kernel
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_gpu.log
timeout 600 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"; cat gpurun_out/bench.json; tail -3 gpurun_out/bench.err
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_reference.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"; cat gpurun_out/bench_reference.json
CMD="python bench.py --steps 1 --warmup 1 --spp 8 --no-cpu-baseline"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_launch.log 2>&1; echo "ncu launch rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_extend_wide -s 3 -c 2 -o gpurun_out/k_extend_wide_r1i -f $CMD > gpurun_out/ncu_full.log 2>&1; echo "ncu full rc=$?"
