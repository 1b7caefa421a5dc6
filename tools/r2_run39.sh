set -x
python -m pytest tests/test_gpu_spec.py -m gpu -x -q -k "saturate" > gpurun_out/r2_s39_pytest_sat.log 2>&1; tail -5 gpurun_out/r2_s39_pytest_sat.log
for w in cfg2 cfg3; do for t in literal structure interp; do python tools/interp_timing.py --workload $w --tier $t --frames 6 2>&1 | tail -1; done; done > gpurun_out/r2_s39_timing.log 2>&1
python - <<'PY'
import json
for l in open('gpurun_out/r2_s39_timing.log'):
    if l.startswith('{'):
        d=json.loads(l); print(d['workload'],d['tier'],round(d['ms'],4),round(d['frac'],4),d['sum'])
PY
python -m pytest tests -m gpu -x -q > gpurun_out/r2_s39_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_s39_pytest_gpu.log
tail -4 gpurun_out/r2_s39_pytest_gpu.log
