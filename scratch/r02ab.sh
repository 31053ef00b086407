#!/bin/bash
# round 2, call ab (1 GPU): why did the default bench line report 2.14 ms for inverse_search when the A/B runs report 1.25?
for fl in "--no-e2e --no-cpu-baseline --no-parity" "--no-cpu-baseline --no-parity" "--no-e2e --no-cpu-baseline --no-parity --steps 20"; do
  echo -n "[$fl] "
  timeout 300 python bench.py --warmup 3 --steps 10 $fl 2>/dev/null | grep '^{' | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l); print(round(d['value'],1), round(d['ms_per_step'],3), d['steps'], d['roofline']['stage_ms_per_step'])"
done
