set -x
T="python tools/interp_timing.py --frames 6"
show() { tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['workload'],d['tier'],'minb',d['minb'],round(d['ms'],4),round(d['frac'],4),d['sum'])"; }
for w in cfg1 cfg2 cfg3; do for t in literal structure interp; do $T --workload $w --tier $t 2>&1 | show; done; done
for m in 6 7; do RM_SPEC_MINB=$m $T --workload cfg2 --tier literal 2>&1 | show; done
python -m pytest tests -m gpu -x -q > gpurun_out/r2_s46_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_s46_pytest_gpu.log
tail -4 gpurun_out/r2_s46_pytest_gpu.log
