set -x
python -m pytest tests/test_gpu_spec.py -m gpu -x -q > gpurun_out/r2_s37_pytest_spec.log 2>&1; tail -5 gpurun_out/r2_s37_pytest_spec.log
for t in literal structure; do python tools/interp_timing.py --workload cfg3 --tier $t --frames 6 2>&1 | tail -1 | cut -c1-60,180-330; done
RM_SPEC_CONSTBANK=0 python tools/interp_timing.py --workload cfg3 --tier structure --frames 6 2>&1 | tail -1 | cut -c1-60,180-330
python -m pytest tests -m gpu -x -q > gpurun_out/r2_s37_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_s37_pytest_gpu.log
tail -4 gpurun_out/r2_s37_pytest_gpu.log
