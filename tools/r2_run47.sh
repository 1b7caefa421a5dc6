set -x
for w in cfg1 cfg2 cfg3; do for t in literal structure interp; do python tools/interp_timing.py --workload $w --tier $t --frames 6 2>&1 | tail -1; done; done > gpurun_out/r2_s47_timing.log 2>&1
cat gpurun_out/r2_s47_timing.log | cut -c1-60,200-330
python -m pytest tests -m gpu -x -q > gpurun_out/r2_s47_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_s47_pytest_gpu.log
tail -5 gpurun_out/r2_s47_pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2_s47_smoke.log 2>&1; tail -3 gpurun_out/r2_s47_smoke.log
python bench.py --steps 10 --warmup 3 > gpurun_out/r2_s47_bench_n1.json 2> gpurun_out/r2_s47_bench_n1.err; echo "bench rc=$?"
B="python bench.py --steps 2 --warmup 3 --strong off --parity off --no-cpu-baseline"
$B > gpurun_out/r2_s47_plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2_s47_launches_cfg3.csv $B > gpurun_out/r2_s47_ncu_launches.log 2>&1
for t in literal structure; do
python tools/ncu_driver.py --tier $t > gpurun_out/r2_s47_plain_$t.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:'rm_spec' -s 2 -c 1 -f -o gpurun_out/r2_s47_$t python tools/ncu_driver.py --tier $t > gpurun_out/r2_s47_ncu_$t.log 2>&1
done
python tools/ncu_driver.py --tier interp > gpurun_out/r2_s47_plain_interp.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:'k_direct' -s 2 -c 1 -f -o gpurun_out/r2_s47_interp python tools/ncu_driver.py --tier interp > gpurun_out/r2_s47_ncu_interp.log 2>&1
python tools/ncu_driver.py --workload cfg2 --tier literal --frames 6 > gpurun_out/r2_s47_plain_cfg2.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:'rm_spec' -s 3 -c 1 -f -o gpurun_out/r2_s47_cfg2 python tools/ncu_driver.py --workload cfg2 --tier literal --frames 6 > gpurun_out/r2_s47_ncu_cfg2.log 2>&1
ls -la gpurun_out | grep r2_s47
# the other single-GPU workloads under the full contract, the opt-in culled mode, the pipelined e2e, the reference arm: final code
for w in cfg2 cfg1 n4 cfg5; do python bench.py --workload $w --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/r2_s47_bench_$w.json 2> gpurun_out/r2_s47_bench_$w.err; echo "bench $w rc=$?"; done
python bench.py --cull on --steps 10 --warmup 3 --strong off --parity off --no-cpu-baseline > gpurun_out/r2_s47_bench_cfg3_cull.json 2> gpurun_out/r2_s47_bench_cfg3_cull.err; echo rc=$?
python bench.py --e2e pipelined --steps 10 --warmup 3 --strong off --parity off --no-cpu-baseline > gpurun_out/r2_s47_bench_cfg3_pipelined.json 2> gpurun_out/r2_s47_pipelined.err; echo rc=$?
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r2_s47_bench_reference.json 2> gpurun_out/r2_s47_bench_reference.err; echo rc=$?
for f in gpurun_out/r2_s47_bench_*.json; do echo $f; cut -c1-300 $f; done
RM_SPEC_R=4 python tools/interp_timing.py --workload cfg3 --tier literal --frames 4 2>&1 | tail -1 | cut -c1-60,200-330
RM_SPEC_R=4 RM_SPEC_NOINLINE=1 python tools/interp_timing.py --workload cfg3 --tier literal --frames 4 2>&1 | tail -1 | cut -c1-60,200-330
