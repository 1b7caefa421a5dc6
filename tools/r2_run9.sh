set -x
O=gpurun_out/r2_s9_crossover.txt
for ni in 0 1; do for r in 2 4; do echo "== NOINLINE=$ni R=$r" >> $O; RM_SPEC_NOINLINE=$ni RM_SPEC_R=$r python tools/spec_crossover.py 8 16 32 48 64 96 128 >> $O 2>&1; done; done
echo "== NOINLINE=1 R=2 MINB=10" >> $O; RM_SPEC_MINB=10 RM_SPEC_NOINLINE=1 RM_SPEC_R=2 python tools/spec_crossover.py 32 64 128 >> $O 2>&1
cat $O
# default settings, the three tiers on cfg3/cfg2 with the current in-tree build
for t in literal structure interp; do for w in cfg3 cfg2; do python tools/interp_timing.py --tier $t --workload $w >> gpurun_out/r2_s9_default.jsonl; done; done
cut -c1-200 gpurun_out/r2_s9_default.jsonl
