# the other single-GPU workloads under the full contract, the opt-in culled mode, the pipelined e2e, the reference arm: final code
set -x
for w in cfg2 cfg1 n4 cfg5; do python bench.py --workload $w --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/r2_s24_bench_$w.json 2> gpurun_out/r2_s24_bench_$w.err; echo "bench $w rc=$?"; done
python bench.py --cull on --steps 10 --warmup 3 --strong off --parity off --no-cpu-baseline > gpurun_out/r2_s24_bench_cfg3_cull.json 2> gpurun_out/r2_s24_bench_cfg3_cull.err; echo rc=$?
python bench.py --e2e pipelined --steps 10 --warmup 3 --strong off --parity off --no-cpu-baseline > gpurun_out/r2_s24_bench_cfg3_pipelined.json 2> gpurun_out/r2_s24_pipelined.err; echo rc=$?
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r2_s24_bench_reference.json 2> gpurun_out/r2_s24_bench_reference.err; echo rc=$?
for f in gpurun_out/r2_s24_bench_*.json; do echo $f; cut -c1-300 $f; done
