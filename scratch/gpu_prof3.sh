#!/bin/bash
# ncu --set full of the level-3 kernels (+ upsample) of one step at batch 592, split TV path
OUT=gpurun_out/${1:-r01f}; mkdir -p $OUT
export DISB200_TV_PATH=split
timeout 500 ncu --set full --clock-control none --import-source on -s $((3*68+49)) -c 19 -o $OUT/l3_full -f \
  python bench.py --steps 2 --warmup 3 --batch 592 --no-e2e --no-cpu-baseline > $OUT/ncu_full.log 2>&1; echo "ncu full rc=$?"; tail -2 $OUT/ncu_full.log
