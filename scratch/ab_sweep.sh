#!/bin/bash
# one process per variant; results collected in gpurun_out/ab.log
out=gpurun_out/ab.log; : > $out
for fmt in bf16 q4; do
  python scratch/ab_decode.py $fmt >> $out 2>&1
  B200_NO_ATTN_CLUSTER=1 python scratch/ab_decode.py $fmt >> $out 2>&1
  B200_TRACE=gpurun_out/trace_$fmt.txt python scratch/ab_decode.py $fmt 16 >> $out 2>&1
done
grep -E "^AB|Error|error" $out
for fmt in bf16 q4; do python scratch/trace_report.py gpurun_out/trace_$fmt.txt; done
