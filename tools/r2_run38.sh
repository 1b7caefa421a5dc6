set -x
python bench.py --steps 10 --warmup 3 > gpurun_out/r2_s38_bench_n1.json 2> gpurun_out/r2_s38_bench_n1.err; echo "bench rc=$?"
python tools/ncu_driver.py --tier structure > gpurun_out/r2_s38_plain_structure.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:'rm_spec' -s 2 -c 1 -f -o gpurun_out/r2_s38_structure python tools/ncu_driver.py --tier structure > gpurun_out/r2_s38_ncu_structure.log 2>&1
tail -2 gpurun_out/r2_s38_plain_structure.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2_s38_smoke.log 2>&1; tail -2 gpurun_out/r2_s38_smoke.log
