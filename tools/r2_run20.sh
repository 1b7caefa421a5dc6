set -x
python bench.py --steps 5 --warmup 3 --strong off --parity off --no-cpu-baseline > gpurun_out/r2_s20_bench_quick.json 2> gpurun_out/r2_s20_bench_quick.err; echo rc=$?
python tools/ncu_driver.py --workload cfg2 --tier literal --frames 6 > gpurun_out/r2_s20_plain_cfg2.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:'rm_spec' -s 3 -c 1 -f -o gpurun_out/r2_s20_cfg2 python tools/ncu_driver.py --workload cfg2 --tier literal --frames 6 > gpurun_out/r2_s20_ncu_cfg2.log 2>&1
tail -2 gpurun_out/r2_s20_plain_cfg2.log
