#!/bin/bash
# scripts/gpu.sh <job> [tag] — the ONE parameterised recipe for `gpurun -- 'bash scripts/gpu.sh <job> <tag>'`.
# Every job writes into gpurun_out/<tag>_*; bench values are never taken under ncu (the plain run
# comes first, ncu only after it exited 0).  Jobs:
#   check     smoke + pytest -m gpu + default bench (both arms)
#   bench     default bench only (+ any extra args after the tag)
#   launches  plain bench, then the ncu launch list of the same command
#   full K    plain bench, then ncu --set full of kernel regex K (2 launches)
#   localmem  plain bench, then the local-memory / LSU counters of k_extend_wide
#   buildncu  plain bench, then ncu --set full of every build kernel (one launch each)
#   c5        C5 workload bench
#   configs   BASELINE configs 2 and 3 at their stated size through the headless driver (needs data/ref_files/,
#             scripts/pack_ref_data.py) + their parity tests
#   sweep     a list of --opt settings given after the tag, one short bench each
#   ab        A/B of library builds: each remaining argument is "LIBSUFFIX|bench args" (LIBSUFFIX "" = librtb200.so,
#             "cs" = librtb200_cs.so built with make -C csrc OUT=../librtb200_cs.so EXTRA=-D...), one short bench each
#   libtest   SUFFIX: the GPU parity tests against librtb200_SUFFIX.so
#   cubsort   library yardstick for the sort stage: cub::DeviceRadixSort::SortPairs on 10 M and 99.7 M (key, id) pairs
#             (scripts/yardstick/cub_sort.cu, scratch binary in build/, never part of the product)
set -u
JOB=${1:-check}; TAG=${2:-run}; shift 2 2>/dev/null || true
O=gpurun_out; mkdir -p $O
SHORT="python bench.py --steps 1 --warmup 1 --spp 8 --no-cpu-baseline --no-c5"
brief() { python - "$1" <<'EOF'
import json, sys
for l in open(sys.argv[1]):
    if l.startswith('{'):
        d = json.loads(l); r = d.get('roofline') or {}; b = d.get('build') or {}
        print('value %.1f e2e %.1f ms/step %.2f trace_ms %.2f lanes? build_ms %.3f (%.0f Mprims/s) launches %s clocks %s' % (
            d['value'], d['e2e']['value'], d['ms_per_step'], r.get('trace_ms_per_step') or 0, b.get('ms') or 0,
            b.get('mprims_per_s') or 0, d.get('gpu_launches'), (d.get('clocks') or {}).get('sm_mhz')))
EOF
}
case $JOB in
cubsort)
  mkdir -p build
  nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a scripts/yardstick/cub_sort.cu -o build/cub_sort && \
  (build/cub_sort 10000010; build/cub_sort 99700000) > $O/${TAG}_cub_sort.txt 2>&1; echo "rc=$?"; cat $O/${TAG}_cub_sort.txt
  ;;
check)
  nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $O/${TAG}_smi.txt
  timeout 300 python __graft_entry__.py smoke > $O/${TAG}_smoke.log 2>&1; echo "smoke rc=$?"; tail -1 $O/${TAG}_smoke.log
  timeout 1500 python -m pytest tests -m gpu -x -q > $O/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 $O/${TAG}_pytest.log
  timeout 900 python bench.py "$@" > $O/${TAG}_bench.json 2> $O/${TAG}_bench.err; echo "bench rc=$?"; brief $O/${TAG}_bench.json; tail -2 $O/${TAG}_bench.err
  timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > $O/${TAG}_bench_reference.json 2> $O/${TAG}_bench_reference.err; echo "ref rc=$?"; cut -c1-300 $O/${TAG}_bench_reference.json
  ;;
bench)
  timeout 900 python bench.py "$@" > $O/${TAG}_bench.json 2> $O/${TAG}_bench.err; echo "bench rc=$?"; brief $O/${TAG}_bench.json; tail -2 $O/${TAG}_bench.err
  ;;
launches)
  CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-c5 $*"
  timeout 600 $CMD > $O/${TAG}_bench.json 2> $O/${TAG}_bench.err && \
  timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $O/${TAG}_launches.csv $CMD > $O/${TAG}_ncu.log 2>&1
  echo "rc=$?"; brief $O/${TAG}_bench.json
  ;;
full)
  K=$1; shift
  CMD="$SHORT $*"
  timeout 600 $CMD > $O/${TAG}_plain.log 2>&1 && \
  timeout 1200 ncu --set full --clock-control none --import-source on -k regex:$K -s ${SKIP:-1} -c ${COUNT:-2} -o $O/${TAG}_full -f $CMD > $O/${TAG}_ncu.log 2>&1
  echo "rc=$?"; tail -2 $O/${TAG}_ncu.log
  ;;
localmem)
  CMD="$SHORT $*"
  M=gpu__time_duration.sum,smsp__inst_executed.sum,smsp__thread_inst_executed_per_inst_executed.ratio,smsp__inst_executed_op_local_ld.sum,smsp__inst_executed_op_local_st.sum,smsp__inst_executed_op_shared_ld.sum,smsp__inst_executed_op_shared_st.sum,smsp__inst_executed_op_global_ld.sum,l1tex__t_bytes_pipe_lsu_mem_local_op_ld.sum,l1tex__t_bytes_pipe_lsu_mem_local_op_st.sum,l1tex__t_bytes_pipe_lsu_mem_global_op_ld.sum,l1tex__t_bytes.sum,l1tex__throughput.avg.pct_of_peak_sustained_elapsed,l1tex__data_pipe_lsu_wavefronts_mem_shared.sum,l1tex__data_pipe_lsu_wavefronts.sum,smsp__issue_active.avg.pct_of_peak_sustained_active,sm__warps_active.avg.pct_of_peak_sustained_active,launch__registers_per_thread,lts__t_sector_hit_rate.pct,l1tex__t_sector_hit_rate.pct,dram__bytes_read.sum,dram__bytes_write.sum
  timeout 600 $CMD > $O/${TAG}_plain.log 2>&1 && \
  timeout 900 ncu --metrics $M --clock-control none -k regex:k_extend_wide -c 6 --csv --log-file $O/${TAG}_localmem.csv $CMD > $O/${TAG}_ncu.log 2>&1
  echo "rc=$?"; tail -2 $O/${TAG}_ncu.log
  ;;
buildncu)
  CMD="python bench.py --steps 1 --warmup 1 --spp 1 --no-cpu-baseline $*"
  timeout 600 $CMD > $O/${TAG}_plain.log 2>&1 && \
  timeout 1500 ncu --set full --clock-control none --import-source on -k regex:'k_(flatten|morton|sort_hist|sort_pass|gather|hierarchy|refit|collapse|wide)' -s 14 -c 14 -o $O/${TAG}_build -f $CMD > $O/${TAG}_ncu.log 2>&1
  echo "rc=$?"; tail -2 $O/${TAG}_ncu.log
  ;;
c5)
  timeout 1200 python bench.py --workload c5 --steps 4 --warmup 3 "$@" > $O/${TAG}_c5.json 2> $O/${TAG}_c5.err; echo "c5 rc=$?"; brief $O/${TAG}_c5.json; tail -2 $O/${TAG}_c5.err
  ;;
configs)
  D=nvidia-optix-ray-tracer_b200/rtb200_render
  timeout 900 python -m pytest tests/test_gpu_configs.py -m gpu -x -q > $O/${TAG}_configs_pytest.log 2>&1; echo "pytest rc=$?"; tail -2 $O/${TAG}_configs_pytest.log
  # C2: particle.vtk.series + wall.vtk.series, all 401 files (sub-frame 0 of each), 1080p, 16 spp, depth 8, rebuild per frame;
  # the camera-facing wall is shape 2 after the reference's transform chain (tests/test_gpu_configs.py)
  timeout 900 $D data/ref_files/config_c2.json --width 1920 --height 1080 --spp 16 --depth 8 --extra-series wall.vtk.series \
      --drop-shape 2 --frame-stride 9 --format none --quiet "$@" > $O/${TAG}_c2.json 2> $O/${TAG}_c2.err; echo "c2 rc=$?"; cat $O/${TAG}_c2.json
  # the same with one PPM per frame written (file output inside the loop)
  timeout 900 $D data/ref_files/config_c2.json --width 1920 --height 1080 --spp 16 --depth 8 --extra-series wall.vtk.series \
      --drop-shape 2 --frame-stride 9 --file-stride 40 --format ppm --out /tmp/${TAG}_c2_frames --quiet "$@" > $O/${TAG}_c2_ppm.json 2>> $O/${TAG}_c2.err; echo "c2 ppm rc=$?"
  python scripts/ppm2png.py /tmp/${TAG}_c2_frames/frame_00400_000.ppm $O/${TAG}_c2_frame400.png 2
  # C3: Mesh mode; "cache": true writes the cache files and exits (Main.cu:23-25), the second run renders from them: 4K, 64 spp, depth 5
  sed 's/"cache": false/"cache": true/' data/ref_files/config_c3.json > data/ref_files/config_c3_cache.json
  rm -rf data/ref_files/cache; T0=$(date +%s%N)
  timeout 900 $D data/ref_files/config_c3_cache.json 2> $O/${TAG}_c3.err; echo "c3 cache rc=$? in $(( ($(date +%s%N) - T0) / 1000000 )) ms, $(ls data/ref_files/cache | wc -l) files" | tee $O/${TAG}_c3_cache.txt; tail -1 $O/${TAG}_c3.err
  timeout 900 $D data/ref_files/config_c3.json --width 3840 --height 2160 --spp 64 --depth 5 --frame-stride 9 --format none --quiet "$@" \
      > $O/${TAG}_c3.json 2>> $O/${TAG}_c3.err; echo "c3 rc=$?"; cat $O/${TAG}_c3.json
  timeout 300 $D data/ref_files/config_c3.json --width 3840 --height 2160 --spp 64 --depth 5 --frame-stride 9 --file-stride 70 --format ppm \
      --out /tmp/${TAG}_c3_frames --quiet "$@" > /dev/null 2>> $O/${TAG}_c3.err && python scripts/ppm2png.py /tmp/${TAG}_c3_frames/frame_00070_000.ppm $O/${TAG}_c3_frame70.png 4
  ;;
sweep)
  # each remaining argument is one quoted option string, e.g. "--opt t1=8 --opt t2=8"
  i=0
  for o in "$@"; do
    i=$((i+1))
    timeout 600 python bench.py --steps 2 --warmup 2 --no-cpu-baseline $o > $O/${TAG}_s$i.json 2> $O/${TAG}_s$i.err
    echo "== [$i] $o rc=$?"; brief $O/${TAG}_s$i.json
  done
  ;;
ab)
  i=0
  for item in "$@"; do
    i=$((i+1)); suf=${item%%|*}; args=${item#*|}
    lib=$PWD/nvidia-optix-ray-tracer_b200/librtb200${suf:+_$suf}.so
    RTB200_LIB=$lib timeout 600 python bench.py --steps 3 --warmup 2 --no-cpu-baseline --no-build-100m $args > $O/${TAG}_ab$i.json 2> $O/${TAG}_ab$i.err
    echo "== [$i] lib=${suf:-default} $args rc=$?"; brief $O/${TAG}_ab$i.json
  done
  ;;
libtest)
  suf=$1
  RTB200_LIB=$PWD/nvidia-optix-ray-tracer_b200/librtb200_$suf.so timeout 1200 python -m pytest tests/test_gpu_parity.py tests/test_gpu_fullsize.py tests/test_gpu_fused.py tests/test_gpu_build.py tests/test_gpu_refparity.py -m gpu -x -q > $O/${TAG}_libtest_$suf.log 2>&1
  echo "libtest $suf rc=$?"; tail -3 $O/${TAG}_libtest_$suf.log
  ;;
*) echo "unknown job $JOB"; exit 2;;
esac
