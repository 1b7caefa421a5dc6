set -x
for w in cfg1 cfg2 cfg3; do for t in literal structure interp; do python tools/interp_timing.py --workload $w --tier $t --frames 6 2>&1 | tail -1; done; done > gpurun_out/r2_s29_timing.log 2>&1
python - <<'PY'
import json
for l in open('gpurun_out/r2_s29_timing.log'):
    if l.startswith('{'):
        d=json.loads(l); print(d['workload'],d['tier'],round(d['ms'],4),round(d['frac'],4))
PY
python -m pytest tests -m gpu -x -q > gpurun_out/r2_s29_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_s29_pytest_gpu.log
tail -4 gpurun_out/r2_s29_pytest_gpu.log
