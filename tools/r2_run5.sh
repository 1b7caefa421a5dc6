set -x
L=variants/librm_p.so
O=gpurun_out/r2_s5_matrix.jsonl
for f in 0 1; do for t in literal structure; do RM_SPEC_FUSED=$f python tools/interp_timing.py --lib $L --tier $t --workload cfg3 >> $O 2>> gpurun_out/r2_s5.err; done; done
for mb in 10 11 13 14; do RM_SPEC_MINB=$mb RM_SPEC_FUSED=1 python tools/interp_timing.py --lib $L --tier literal --workload cfg3 >> $O 2>> gpurun_out/r2_s5.err; done
for v in "" 8o 8t; do RM_INTERP_VARIANT=$v python tools/interp_timing.py --lib $L --workload cfg3 >> $O 2>> gpurun_out/r2_s5.err; done
for v in "" 8o 8t; do RM_INTERP_VARIANT=$v python tools/interp_timing.py --lib $L --workload cfg2 >> $O 2>> gpurun_out/r2_s5.err; done
for t in literal structure; do python tools/interp_timing.py --lib $L --tier $t --workload cfg2 >> $O 2>> gpurun_out/r2_s5.err; done
cat $O; tail -5 gpurun_out/r2_s5.err
