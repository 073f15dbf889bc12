mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_build.py tests/test_gpu_parity.py -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/pytest_gpu.log
