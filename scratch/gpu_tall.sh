#!/bin/bash
run() { echo "== $*"; timeout 300 python bench.py --steps 4 --warmup 3 --no-e2e --no-cpu-baseline "$@" 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print(round(d['value'],1), round(d['ms_per_step'],3), d['gpu_launches'], d['roofline']['stage_ms_per_step'])
    else: print(l.rstrip()[-300:])"; }
timeout 600 python -m pytest tests/test_parity_gpu.py -m gpu -x -q -k "tv_paths or tall_level or c4 or c2" 2>&1 | tail -3
run --workload C4 --batch 1
run --workload C4 --batch 16
DISB200_SORB_MAX=100000 run --workload C4 --batch 64
DISB200_SORB_MAX=0 run --workload C4 --batch 64
DISB200_SORB_MAX=100000 run --workload op3 --batch 64
DISB200_SORB_MAX=0 run --workload op3 --batch 64
