set -x
python -m pytest tests -m gpu -x -q > gpurun_out/r2_s48_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_s48_pytest_gpu.log
tail -3 gpurun_out/r2_s48_pytest_gpu.log
python bench.py --steps 10 --warmup 3 > gpurun_out/r2_s48_bench_n1.json 2> gpurun_out/r2_s48_bench_n1.err; echo "bench rc=$?"
for w in cfg2 cfg5; do python bench.py --workload $w --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/r2_s48_bench_$w.json 2> gpurun_out/r2_s48_bench_$w.err; echo "bench $w rc=$?"; done
B="python bench.py --steps 2 --warmup 3 --strong off --parity off --no-cpu-baseline"
$B > gpurun_out/r2_s48_plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2_s48_launches_cfg3.csv $B > gpurun_out/r2_s48_ncu_launches.log 2>&1
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2_s48_smoke.log 2>&1; tail -1 gpurun_out/r2_s48_smoke.log
