set -x
python tools/interp_timing.py --workload cfg3 --tier structure --frames 6 2>&1 | tail -1 | cut -c1-60,180-330
RM_SPEC_NOINLINE=1 python tools/interp_timing.py --workload cfg3 --tier structure --frames 6 2>&1 | tail -1 | cut -c1-60,180-330
RM_SPEC_NOINLINE=1 RM_SPEC_CONSTBANK=1 python tools/interp_timing.py --workload cfg3 --tier structure --frames 6 2>&1 | tail -1 | cut -c1-60,180-330
for m in 8 9 11 12; do RM_SPEC_MINB=$m RM_SPEC_NOINLINE=1 RM_SPEC_CONSTBANK=1 python tools/interp_timing.py --workload cfg3 --tier structure --frames 6 2>&1 | tail -1 | cut -c1-60,180-330; done
