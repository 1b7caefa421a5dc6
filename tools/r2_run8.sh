set -x
L=variants/librm_n.so
O=gpurun_out/r2_s8_noinline_sweep.jsonl
for mb in 7 8 9 10 11; do RM_SPEC_MINB=$mb RM_SPEC_NOINLINE=1 python tools/interp_timing.py --lib $L --tier literal --workload cfg3 >> $O 2>> gpurun_out/r2_s8.err; done
for mb in 9 10 11; do RM_SPEC_MINB=$mb RM_SPEC_NOINLINE=0 python tools/interp_timing.py --lib $L --tier literal --workload cfg3 >> $O 2>> gpurun_out/r2_s8.err; done
for mb in 8 9 10 11; do RM_SPEC_MINB=$mb RM_SPEC_NOINLINE=1 python tools/interp_timing.py --lib $L --tier structure --workload cfg3 >> $O 2>> gpurun_out/r2_s8.err; done
for mb in 6 8 10 12; do RM_SPEC_MINB=$mb RM_SPEC_NOINLINE=1 python tools/interp_timing.py --lib $L --tier literal --workload cfg2 >> $O 2>> gpurun_out/r2_s8.err; done
for mb in 8 12; do RM_SPEC_MINB=$mb RM_SPEC_NOINLINE=0 python tools/interp_timing.py --lib $L --tier literal --workload cfg2 >> $O 2>> gpurun_out/r2_s8.err; done
for r in 2 8; do RM_SPEC_R=$r RM_SPEC_NOINLINE=1 python tools/interp_timing.py --lib $L --tier literal --workload cfg2 >> $O 2>> gpurun_out/r2_s8.err; done
for mb in 8 10 12; do RM_SPEC_MINB=$mb RM_SPEC_NOINLINE=1 python tools/interp_timing.py --lib $L --tier literal --workload cfg1 >> $O 2>> gpurun_out/r2_s8.err; done
RM_SPEC_NOINLINE=0 python tools/interp_timing.py --lib $L --tier literal --workload cfg1 >> $O 2>> gpurun_out/r2_s8.err
for mb in 8 12; do RM_SPEC_MINB=$mb RM_SPEC_NOINLINE=1 python tools/interp_timing.py --lib $L --tier literal --workload n4 >> $O 2>> gpurun_out/r2_s8.err; done
cat $O | cut -c1-300; tail -5 gpurun_out/r2_s8.err
