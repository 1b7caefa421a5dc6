set -x
L=variants/librm_n.so
O=gpurun_out/r2_s7_noinline.jsonl
for ni in 0 1; do for t in literal structure; do for w in cfg3 cfg2; do RM_SPEC_NOINLINE=$ni python tools/interp_timing.py --lib $L --tier $t --workload $w >> $O 2>> gpurun_out/r2_s7.err; done; done; done
for mb in 10 14 16; do RM_SPEC_MINB=$mb RM_SPEC_NOINLINE=1 python tools/interp_timing.py --lib $L --tier literal --workload cfg3 >> $O 2>> gpurun_out/r2_s7.err; done
for r in 4; do RM_SPEC_R=$r RM_SPEC_NOINLINE=1 python tools/interp_timing.py --lib $L --tier literal --workload cfg3 >> $O 2>> gpurun_out/r2_s7.err; done
RM_SPEC_NOINLINE=1 python tools/interp_timing.py --lib $L --tier literal --workload n4 >> $O 2>> gpurun_out/r2_s7.err
RM_SPEC_NOINLINE=0 python tools/interp_timing.py --lib $L --tier literal --workload n4 >> $O 2>> gpurun_out/r2_s7.err
cat $O; tail -5 gpurun_out/r2_s7.err
