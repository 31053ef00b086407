#!/bin/bash
# round 2, call w (1 GPU): ncu --set full of the SOR launches of one step (batch 592) after the team schedule became the default, to
# replace the SOR rows of the r02s whole-step capture; details page of the level-3 team launches at the bench batch
O=gpurun_out
T=/tmp/ncu_r02w; mkdir -p $T
B="python bench.py --steps 2 --warmup 3 --streams 1 --no-e2e --no-cpu-baseline --no-parity"
$B --batch 592 > $O/r02w_plain_592.log 2>&1 && ncu --set full --clock-control none -k regex:"sor_" -s 45 -c 15 -f -o $T/sor_b592 $B --batch 592 > /dev/null 2>&1
echo "sor step rc=$?"
ncu -i $T/sor_b592.ncu-rep --page raw --csv > $O/r02w_sor_full_raw_b592.csv 2>/dev/null; wc -l $O/r02w_sor_full_raw_b592.csv
$B > /dev/null 2>&1 && ncu --set full --clock-control none --import-source on -k regex:"sor_team" -s 15 -c 2 -f -o $T/team_b1776 $B > /dev/null 2>&1
echo "team rc=$?"
ncu -i $T/team_b1776.ncu-rep --page raw --csv > $O/r02w_team_full_raw_b1776.csv 2>/dev/null
ncu -i $T/team_b1776.ncu-rep --page details > $O/r02w_team_details_b1776.txt 2>/dev/null
ncu -i $T/team_b1776.ncu-rep --page source --csv 2>/dev/null | gzip -9 > $O/r02w_team_source_b1776.csv.gz
ls -la $O/ | grep r02w
