#!/bin/bash
# First GPU pass: parity tests, smoke, bench, ncu launch list + one full capture of the traversal kernel.
set -x
mkdir -p gpurun_out
nvidia-smi > gpurun_out/nvidia-smi.txt 2>&1
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/smoke.log
tail -3 gpurun_out/smoke.log
timeout 900 python bench.py --steps 2 --warmup 3 > gpurun_out/bench.log 2> gpurun_out/bench.err; echo "bench rc=$?"
cat gpurun_out/bench.log; tail -5 gpurun_out/bench.err
# ncu: launch list of a reduced-spp run of the same command (per-launch times are cold-cache/serialised)
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv \
   python bench.py --steps 1 --warmup 1 --spp 8 --no-cpu-baseline > gpurun_out/ncu_launch.log 2>&1; echo "ncu launch rc=$?"
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:k_bounce -s 2 -c 2 -o gpurun_out/k_bounce_full -f \
   python bench.py --steps 1 --warmup 1 --spp 8 --no-cpu-baseline > gpurun_out/ncu_full.log 2>&1; echo "ncu full rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"k_primary|k_refit|k_sort_scatter|k_hierarchy" -c 8 -o gpurun_out/build_full -f \
   python bench.py --steps 1 --warmup 0 --spp 4 --no-cpu-baseline > gpurun_out/ncu_full2.log 2>&1; echo "ncu full2 rc=$?"
ls -la gpurun_out
