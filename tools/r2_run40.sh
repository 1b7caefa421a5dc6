set -x
python -m pytest tests -m gpu -x -q > gpurun_out/r2_s40_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_s40_pytest_gpu.log
tail -6 gpurun_out/r2_s40_pytest_gpu.log
