#!/bin/bash
# round-2 ncu evidence: launch lists (device time per launch) and full captures of the decode kernels
set -x
O=gpurun_out
for fmt in bf16 q4; do
  python profiles/prof_decode.py $fmt > $O/plain_$fmt.log 2>&1 || exit 1
  ncu --metrics gpu__time_duration.sum --clock-control none --kernel-name-base demangled -k regex:"gemv|attn_decode|finish_step|sample" -c 1000 \
      --csv --log-file $O/launches_r2_decode_$fmt.csv python profiles/prof_decode.py $fmt > $O/ncu_l_$fmt.log 2>&1
done
# full captures: last decode step's kernels (gemv launches: 4 per layer + lm_head; skip 4 steps = 4 x 129)
ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k regex:gemv_kernel -s 520 -c 5 -o $O/prof_gemv_bf16_r2 -f \
    python profiles/prof_decode.py bf16 > $O/ncu_f_bf16.log 2>&1
ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k regex:"gemv_kernel|gemv_dual" -s 520 -c 5 -o $O/prof_gemv_q4_r2 -f \
    python profiles/prof_decode.py q4 > $O/ncu_f_q4.log 2>&1
ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k regex:attn_decode_cluster -s 130 -c 2 -o $O/prof_attn_cluster_r2 -f \
    python profiles/prof_decode.py bf16 > $O/ncu_f_attn.log 2>&1
for n in prof_gemv_bf16_r2 prof_gemv_q4_r2 prof_attn_cluster_r2; do
  ncu -i $O/$n.ncu-rep --page raw --csv > $O/${n}_raw.csv 2>/dev/null
  rm -f $O/$n.ncu-rep
done
ls -la $O/*.csv
