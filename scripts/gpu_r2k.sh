#!/bin/bash
# r2k: occupancy variants of k_sort_pass / k_collapse, rebuilt on the box (launch lists only)
mkdir -p gpurun_out
variant() {
  name=$1; shift
  touch nvidia-optix-ray-tracer_b200/csrc/rt_build.cu
  make -C nvidia-optix-ray-tracer_b200/csrc EXTRA="$*" > gpurun_out/make_$name.log 2>&1 || { echo "$name: build failed"; return; }
  timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/launches_$name.csv python bench.py --steps 1 --warmup 1 --spp 4 --no-cpu-baseline > gpurun_out/ncu_$name.log 2>&1
  echo "== $name ($*)"; python scripts/launch_summary.py gpurun_out/launches_$name.csv | grep -E "collapse|sort_pass"
}
variant c6 -DRT_COLLAPSE_MIN_BLOCKS=6
variant c5 -DRT_COLLAPSE_MIN_BLOCKS=5
variant c4 -DRT_COLLAPSE_MIN_BLOCKS=4
variant c3 -DRT_COLLAPSE_MIN_BLOCKS=3
touch nvidia-optix-ray-tracer_b200/csrc/rt_build.cu; make -C nvidia-optix-ray-tracer_b200/csrc > /dev/null 2>&1
