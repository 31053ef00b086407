#!/bin/bash
OUT=gpurun_out/sorw2; mkdir -p $OUT
DISB200_TV_PATH=split timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -5
for cfg in "256 fused" "256 split" "1480 split" "2960 split"; do set -- $cfg
  echo "== batch $1 path $2"; DISB200_TV_PATH=$2 timeout 300 python bench.py --steps 10 --warmup 3 --batch $1 --no-e2e --no-cpu-baseline 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print(d['value'], d['ms_per_step'], d['gpu_launches'], d['roofline']['stage_ms_per_step'])
    else: print(l.rstrip())"
done
