mkdir -p gpurun_out
CMD="python bench.py --steps 1 --warmup 1 --spp 4 --no-cpu-baseline"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'k_refit' -c 1 -o gpurun_out/refit_r1m -f $CMD > gpurun_out/ncu_full.log 2>&1; echo "ncu full rc=$?"
