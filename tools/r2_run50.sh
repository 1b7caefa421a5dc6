python bench.py --steps 3 --warmup 3 --strong off --parity off --no-cpu-baseline > gpurun_out/r2_s50_bench_quick.json 2> gpurun_out/r2_s50_bench_quick.err; echo "rc=$?"
