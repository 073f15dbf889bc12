#!/bin/bash
# r1p: rt_spheres_advect tests + C5 workload (100M-sphere DEM packing, frame-parallel) on one GPU
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_advect.py -m gpu -x -q > gpurun_out/pytest_advect.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_advect.log
timeout 900 python bench.py --workload c5 --n-spheres 20000000 --steps 2 --warmup 1 > gpurun_out/bench_c5_20m.json 2> gpurun_out/bench_c5_20m.err; echo "c5 20M rc=$?"; tail -c 2500 gpurun_out/bench_c5_20m.json; tail -3 gpurun_out/bench_c5_20m.err
timeout 1200 python bench.py --workload c5 --steps 3 --warmup 3 > gpurun_out/bench_c5.json 2> gpurun_out/bench_c5.err; echo "c5 100M rc=$?"; tail -c 2500 gpurun_out/bench_c5.json; tail -3 gpurun_out/bench_c5.err
