#!/bin/bash
run() { echo "== $*"; timeout 300 python bench.py --steps 10 --warmup 3 --no-e2e --no-cpu-baseline "$@" 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print(round(d['value'],1), round(d['ms_per_step'],3), d['gpu_launches'])
    else: print(l.rstrip()[-300:])"; }
run --batch 1480
DISB200_STREAMS=2 run --batch 1480
DISB200_STREAMS=2 run --batch 2960
DISB200_STREAMS=3 run --batch 4440
DISB200_STREAMS=4 run --batch 2960
