set -x
python -m pytest tests/test_gpu_cull.py -m gpu -q > gpurun_out/r2_s12_pytest.log 2>&1; tail -12 gpurun_out/r2_s12_pytest.log
python bench.py --cull on --steps 10 --warmup 3 --strong off --parity off > gpurun_out/r2_s12_bench_cfg3_cull.json 2> gpurun_out/r2_s12_bench_cfg3_cull.err; echo rc=$?; tail -3 gpurun_out/r2_s12_bench_cfg3_cull.err
python bench.py --cull on --workload cfg2 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/r2_s12_bench_cfg2_cull.json 2>&1
