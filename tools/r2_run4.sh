# interpreter measurement variants (profiles/r2_interp_NOTES.md)
set -x
for lib in variants/librm_x4.so variants/librm_x3.so; do
  for v in "" 8t 16 16t; do
    RM_INTERP_VARIANT=$v python tools/interp_timing.py --lib $lib --workload cfg3 >> gpurun_out/r2_s4_interp.jsonl 2>> gpurun_out/r2_s4_interp.err
    RM_INTERP_VARIANT=$v python tools/interp_timing.py --lib $lib --workload cfg2 >> gpurun_out/r2_s4_interp.jsonl 2>> gpurun_out/r2_s4_interp.err
  done
done
cat gpurun_out/r2_s4_interp.jsonl
tail -5 gpurun_out/r2_s4_interp.err
