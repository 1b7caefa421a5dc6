set -x
python -m pytest tests -m gpu -x -q > gpurun_out/r2_s1_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_s1_pytest_gpu.log
tail -5 gpurun_out/r2_s1_pytest_gpu.log
python bench.py --steps 10 --warmup 3 > gpurun_out/r2_s1_bench_n1.json 2> gpurun_out/r2_s1_bench_n1.err; echo "bench rc=$?"
tail -c 600 gpurun_out/r2_s1_bench_n1.err
python tools/n4_timing.py > gpurun_out/r2_s1_n4_plain.json 2>&1
RM_N4_COLLAPSE=1 python tools/n4_timing.py --variants interpreter > gpurun_out/r2_s1_n4_collapse.json 2>&1
RM_N4_COLLAPSE=1 python -m pytest tests/test_n4_gpu.py -m gpu -q > gpurun_out/r2_s1_n4_collapse_pytest.log 2>&1; tail -3 gpurun_out/r2_s1_n4_collapse_pytest.log
cat gpurun_out/r2_s1_n4_plain.json gpurun_out/r2_s1_n4_collapse.json | cut -c1-400
