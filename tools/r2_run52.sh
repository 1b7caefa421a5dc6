python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2_s52_smoke.log 2>&1; tail -2 gpurun_out/r2_s52_smoke.log
python tools/interp_timing.py --workload cfg3 --tier literal --frames 3 2>&1 | tail -1 | cut -c1-60,200-330
