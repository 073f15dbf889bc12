#!/bin/bash
# GPU pass A: parity tests, smoke, bench (both arms), ncu launch list + one full capture of the traversal kernel.
mkdir -p gpurun_out
nvidia-smi > gpurun_out/nvidia-smi.txt 2>&1
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" | tee -a gpurun_out/smoke.log
tail -3 gpurun_out/smoke.log
timeout 900 python bench.py --steps 2 --warmup 3 > gpurun_out/bench.log 2> gpurun_out/bench.err; echo "bench rc=$?"
cat gpurun_out/bench.log; tail -5 gpurun_out/bench.err
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.log 2> gpurun_out/bench_ref.err; echo "bench ref rc=$?"
cat gpurun_out/bench_ref.log; tail -5 gpurun_out/bench_ref.err
CMD="python bench.py --steps 1 --warmup 1 --spp 8 --no-cpu-baseline"
timeout 600 $CMD > gpurun_out/plain.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_launch.log 2>&1; echo "ncu launch rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_extend_wide -s 3 -c 2 -o gpurun_out/k_extend_wide_full -f $CMD > gpurun_out/ncu_full.log 2>&1; echo "ncu full rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"k_flatten|k_morton|k_sort_hist|k_sort_scatter|k_gather|k_hierarchy|k_refit|k_collapse|k_shade|k_accumulate" -c 24 -o gpurun_out/build_full -f \
   python bench.py --steps 1 --warmup 0 --spp 4 --no-cpu-baseline > gpurun_out/ncu_full2.log 2>&1; echo "ncu full2 rc=$?"
ls -la gpurun_out
