#!/bin/bash
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/c4_launches.csv python bench.py --workload C4 --batch 1 --steps 1 --warmup 1 --no-e2e --no-cpu-baseline > gpurun_out/c4_ncu.log 2>&1
tail -2 gpurun_out/c4_ncu.log | cut -c1-300
