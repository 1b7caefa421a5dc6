set -x
# the allocation-free entry point: alone in a fresh process, then at the end of the whole suite under -x (as the round's end runs it)
python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > gpurun_out/r2_s15_pytest_zz_alone.log 2>&1; echo "zz alone rc=$?" >> gpurun_out/r2_s15_pytest_zz_alone.log
tail -6 gpurun_out/r2_s15_pytest_zz_alone.log
python -m pytest tests -m gpu -x -q > gpurun_out/r2_s15_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_s15_pytest_gpu.log
tail -8 gpurun_out/r2_s15_pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2_s15_smoke.log 2>&1; tail -3 gpurun_out/r2_s15_smoke.log
python bench.py --steps 10 --warmup 3 > gpurun_out/r2_s15_bench_n1.json 2> gpurun_out/r2_s15_bench_n1.err; echo "bench rc=$?"
B="python bench.py --steps 2 --warmup 3 --strong off --parity off --no-cpu-baseline"
$B > gpurun_out/r2_s15_plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2_s15_launches_cfg3.csv $B > gpurun_out/r2_s15_ncu_launches.log 2>&1
for t in literal structure; do
python tools/ncu_driver.py --tier $t > gpurun_out/r2_s15_plain_$t.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:'rm_spec' -s 2 -c 1 -f -o gpurun_out/r2_s15_$t python tools/ncu_driver.py --tier $t > gpurun_out/r2_s15_ncu_$t.log 2>&1
done
python tools/ncu_driver.py --tier interp --cull > gpurun_out/r2_s15_plain_cull.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:'k_direct|k_cull' -s 4 -c 2 -f -o gpurun_out/r2_s15_cull python tools/ncu_driver.py --tier interp --cull > gpurun_out/r2_s15_ncu_cull.log 2>&1
ls -la gpurun_out | grep r2_s15
