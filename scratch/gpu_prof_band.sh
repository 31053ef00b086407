#!/bin/bash
OUT=gpurun_out/band; mkdir -p $OUT
export PYTHONPATH=.
timeout 400 ncu --set full --clock-control none --import-source on -k regex:sor_band -c 1 -o $OUT/band_full -f python scratch/sorb_stats.py 544 960 > $OUT/ncu.log 2>&1; echo "rc=$?"; tail -2 $OUT/ncu.log
