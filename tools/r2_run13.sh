set -x
python -m pytest tests -m gpu -q > gpurun_out/r2_s13_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_s13_pytest_gpu.log
tail -8 gpurun_out/r2_s13_pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2_s13_smoke.log 2>&1; tail -3 gpurun_out/r2_s13_smoke.log
python bench.py --steps 10 --warmup 3 > gpurun_out/r2_s13_bench_n1.json 2> gpurun_out/r2_s13_bench_n1.err; echo "bench rc=$?"
B="python bench.py --steps 2 --warmup 3 --strong off --parity off --no-cpu-baseline"
$B > gpurun_out/r2_s13_plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2_s13_launches_cfg3.csv $B > gpurun_out/r2_s13_ncu_launches.log 2>&1
python tools/ncu_driver.py --tier literal > gpurun_out/r2_s13_plain_literal.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:'rm_spec' -s 2 -c 1 -f -o gpurun_out/r2_s13_literal python tools/ncu_driver.py --tier literal > gpurun_out/r2_s13_ncu_literal.log 2>&1
python tools/ncu_driver.py --tier structure > gpurun_out/r2_s13_plain_structure.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:'rm_spec' -s 2 -c 1 -f -o gpurun_out/r2_s13_structure python tools/ncu_driver.py --tier structure > gpurun_out/r2_s13_ncu_structure.log 2>&1
python tools/ncu_driver.py --tier interp --cull > gpurun_out/r2_s13_plain_cull.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:'k_direct|k_cull' -s 4 -c 2 -f -o gpurun_out/r2_s13_cull python tools/ncu_driver.py --tier interp --cull > gpurun_out/r2_s13_ncu_cull.log 2>&1
tail -3 gpurun_out/r2_s13_plain_cull.log
ls -la gpurun_out | grep r2_s13
