set -x
python -m pytest tests -m gpu -q > gpurun_out/r2_s10_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_s10_pytest_gpu.log
tail -15 gpurun_out/r2_s10_pytest_gpu.log
python bench.py --steps 10 --warmup 3 > gpurun_out/r2_s10_bench_n1.json 2> gpurun_out/r2_s10_bench_n1.err; echo "bench rc=$?"
tail -c 800 gpurun_out/r2_s10_bench_n1.err
for w in cfg2 cfg1 n4 cfg5; do python bench.py --workload $w --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/r2_s10_bench_$w.json 2> gpurun_out/r2_s10_bench_$w.err; echo "bench $w rc=$?"; done
python bench.py --workload cfg2 --steps 10 --warmup 3 --no-cpu-baseline --specialize off > gpurun_out/r2_s10_bench_cfg2_interp.json 2>&1
for t in literal structure interp; do
  python tools/ncu_driver.py --tier $t > gpurun_out/r2_s10_plain_$t.log 2>&1 && \
  ncu --set full --clock-control none --import-source on -k regex:'rm_spec|k_direct' -s 2 -c 1 -f -o gpurun_out/r2_s10_$t python tools/ncu_driver.py --tier $t > gpurun_out/r2_s10_ncu_$t.log 2>&1
  cat gpurun_out/r2_s10_plain_$t.log | tail -4
done
ls -la gpurun_out/
