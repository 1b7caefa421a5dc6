set -x
for w in cfg1 cfg2 cfg3; do for t in literal structure interp; do python tools/interp_timing.py --workload $w --tier $t --frames 6 2>&1 | tail -1; done; done > gpurun_out/r2_s28_timing.log 2>&1
cat gpurun_out/r2_s28_timing.log | cut -c1-60,200-330
python -m pytest tests -m gpu -x -q > gpurun_out/r2_s28_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_s28_pytest_gpu.log
tail -5 gpurun_out/r2_s28_pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2_s28_smoke.log 2>&1; tail -3 gpurun_out/r2_s28_smoke.log
python bench.py --steps 10 --warmup 3 > gpurun_out/r2_s28_bench_n1.json 2> gpurun_out/r2_s28_bench_n1.err; echo "bench rc=$?"
B="python bench.py --steps 2 --warmup 3 --strong off --parity off --no-cpu-baseline"
