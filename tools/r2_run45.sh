set -x
T="python tools/interp_timing.py --frames 6"
show() { tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['workload'],d['tier'],'minb',d['minb'],round(d['ms'],4),round(d['frac'],4),d['sum'])"; }
$T --workload cfg3 --tier literal 2>&1 | show
$T --workload cfg3 --tier structure 2>&1 | show
RM_SPEC_PACKED=0 $T --workload cfg3 --tier literal 2>&1 | show
for m in 8 9 11 12; do RM_SPEC_MINB=$m $T --workload cfg3 --tier literal 2>&1 | show; done
$T --workload cfg2 --tier literal 2>&1 | show
RM_SPEC_PACKED=0 $T --workload cfg2 --tier literal 2>&1 | show
for m in 8 9; do RM_SPEC_MINB=$m $T --workload cfg2 --tier literal 2>&1 | show; done
$T --workload cfg1 --tier literal 2>&1 | show
python -m pytest tests/test_gpu_spec.py tests/test_gpu_oracle_full_size.py -m gpu -x -q > gpurun_out/r2_s45_pytest.log 2>&1; tail -4 gpurun_out/r2_s45_pytest.log
