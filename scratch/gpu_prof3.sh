#!/bin/bash
OUT=gpurun_out/r01d; mkdir -p $OUT
export DISB200_TV_PATH=split
timeout 400 ncu --set full --clock-control none --import-source on -s $((3*68+49)) -c 9 -o $OUT/l3_full -f \
  python bench.py --steps 2 --warmup 3 --batch 592 --no-e2e --no-cpu-baseline > $OUT/ncu_full.log 2>&1; echo "ncu full rc=$?"; tail -2 $OUT/ncu_full.log
