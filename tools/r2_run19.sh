set -x
O=gpurun_out/r2_s19_structure_matrix.jsonl
for r in 2 4; do for ni in 0 1; do RM_SPEC_R=$r RM_SPEC_NOINLINE=$ni python tools/interp_timing.py --tier structure --workload cfg3 >> $O 2>> gpurun_out/r2_s19.err; done; done
for mb in 8 9 10 11; do RM_SPEC_MINB=$mb python tools/interp_timing.py --tier structure --workload cfg3 >> $O 2>> gpurun_out/r2_s19.err; done
cut -c1-330 $O; tail -3 gpurun_out/r2_s19.err
