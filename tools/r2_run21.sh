set -x
for w in cfg1 cfg2 cfg3; do for t in literal structure interp; do python tools/interp_timing.py --workload $w --tier $t --frames 6 2>&1 | tail -1; done; done > gpurun_out/r2_s21_timing.log 2>&1
cat gpurun_out/r2_s21_timing.log
python -m pytest tests -m gpu -x -q > gpurun_out/r2_s21_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_s21_pytest_gpu.log
tail -8 gpurun_out/r2_s21_pytest_gpu.log
