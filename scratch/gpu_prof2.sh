#!/bin/bash
OUT=gpurun_out/r01c; mkdir -p $OUT
B=${1:-1480}
L=68
timeout 600 ncu --set full --clock-control none --import-source on -s $((3*L)) -c $L -o $OUT/step_full -f \
  python bench.py --steps 2 --warmup 3 --batch $B --no-e2e --no-cpu-baseline > $OUT/ncu_full.log 2>&1; echo "ncu full rc=$?"; tail -2 $OUT/ncu_full.log
