set -x
O=gpurun_out/r2_s18_recomp.jsonl
for lib in rc0 rc7 rc8 rc9 rc10; do
  python tools/interp_timing.py --lib variants/librm_$lib.so --workload cfg3 >> $O 2>> gpurun_out/r2_s18.err
done
for lib in rc0 rc8; do python tools/interp_timing.py --lib variants/librm_$lib.so --workload n4 >> $O 2>> gpurun_out/r2_s18.err; done
cut -c1-260 $O; tail -3 gpurun_out/r2_s18.err
