#!/bin/bash
# ncu --set full of one whole step at batch 592 (68 launches), current kernels, single stream
OUT=gpurun_out/r01i; mkdir -p $OUT
export DISB200_TV_PATH=split DISB200_STREAMS=1
timeout 1100 ncu --set full --clock-control none --import-source on -s $((3*68)) -c 68 -o $OUT/step_full -f \
  python bench.py --steps 2 --warmup 3 --batch 592 --no-e2e --no-cpu-baseline > $OUT/ncu_full.log 2>&1; echo "ncu full rc=$?"; tail -2 $OUT/ncu_full.log | cut -c1-200
ncu -i $OUT/step_full.ncu-rep --page raw --csv > $OUT/step_full_raw.csv 2>/dev/null; wc -l $OUT/step_full_raw.csv
# SASS pages of the level-3 launches of the three heaviest kernels (ids counted from the launch order of a step)
for k in patch_search system_kernel sor_warp; do
  ncu -i $OUT/step_full.ncu-rep --page source --csv --kernel-name regex:$k > $OUT/source_$k.csv 2>/dev/null
done
ls -la $OUT; rm -f $OUT/step_full.ncu-rep
# launch list of two steps at the default batch
export -n DISB200_TV_PATH
DISB200_STREAMS=1 timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -s 204 -c 136 --csv --log-file $OUT/launches_b1480.csv python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu-baseline > $OUT/ncu_launch.log 2>&1; echo "launch list rc=$?"
rm -f $OUT/step_full.ncu-rep.tmp
