#!/bin/bash
# round 2, call s (1 GPU): ncu evidence of the shipped kernels after the staged 8x8 search - launch list (batch 1776), --set full whole step (batch 592),
# source-level capture of the three hot kernels (batch 1776).  The .ncu-rep files stay on the box (gpurun_out is capped at
# 64 MiB): only the CSV / text exports come back.
O=gpurun_out
T=/tmp/ncu_r02s; mkdir -p $T
B="python bench.py --steps 2 --warmup 3 --streams 1 --tv-schedule 2 --no-e2e --no-cpu-baseline --no-parity"
$B > $O/r02s_plain_1776.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -s 204 -c 136 --csv --log-file $O/r02s_launches_C1_b1776.csv $B > /dev/null 2>&1
echo "launch list rc=$?"; tail -1 $O/r02s_plain_1776.log | cut -c1-200
$B --batch 592 > $O/r02s_plain_592.log 2>&1 && ncu --set full --clock-control none -s 204 -c 68 -f -o $T/step_b592 $B --batch 592 > /dev/null 2>&1
echo "full step rc=$?"
ncu -i $T/step_b592.ncu-rep --page raw --csv > $O/r02s_step_full_raw_b592.csv 2>/dev/null; wc -l $O/r02s_step_full_raw_b592.csv
$B > /dev/null 2>&1 && ncu --set full --clock-control none --import-source on -k regex:"patch_search|sor_warp_kernel|system_kernel|upsample" -s 126 -c 10 -f -o $T/hot_b1776 $B > /dev/null 2>&1
echo "hot kernels rc=$?"
ncu -i $T/hot_b1776.ncu-rep --page raw --csv > $O/r02s_hot_full_raw_b1776.csv 2>/dev/null
ncu -i $T/hot_b1776.ncu-rep --page details > $O/r02s_hot_details_b1776.txt 2>/dev/null
ncu -i $T/hot_b1776.ncu-rep --page source --csv 2>/dev/null | gzip -9 > $O/r02s_hot_source_b1776.csv.gz
ls -la $O/ $T/
du -sh $O
