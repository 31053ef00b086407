#!/bin/bash
mkdir -p gpurun_out
echo "== tests pyramid + rgb"; timeout 600 python -m pytest tests -m gpu -x -q -k "pyramid or rgb or c2" 2>&1 | tail -3
run() { # workload batch steps
  timeout 400 python bench.py --workload $1 --batch $2 --steps $3 --warmup 3 --no-e2e --no-cpu-baseline 2>/dev/null | grep '^{' | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l); print(round(d['value'],1), 'pairs/s', round(d['ms_per_step'],3), 'ms', d['config']['pairs_per_step_per_gpu'], d['roofline']['stage_ms_per_step'])"
}
echo "== rgb2"; run rgb2 1184 6
echo "== C2"; run C2 64 2
echo "== C5 592"; run C5 592 4
echo "== C5 888"; run C5 888 4
echo "== C3 888"; run C3 888 6
