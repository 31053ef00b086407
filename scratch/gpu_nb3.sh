#!/bin/bash
timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
for w in C5 C3; do
for path in old split; do
echo "== $w $path"; DISB200_TV_PATH=$path timeout 300 python bench.py --workload $w --steps 6 --warmup 3 --no-e2e --no-cpu-baseline 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print(round(d['value']), round(d['ms_per_step'],3), d['gpu_launches'], d['roofline']['stage_ms_per_step'])
    else: print(l.rstrip()[-300:])"
done; done
SKIPTEST=1 bash scratch/gpu_quick.sh "1480 split"
