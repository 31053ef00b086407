#!/bin/bash
# usage: gpu_quick.sh "<batch path>" ...   (runs gpu tests first unless SKIPTEST=1)
if [ -z "$SKIPTEST" ]; then DISB200_TV_PATH=split timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3; fi
for cfg in "$@"; do set -- $cfg
  echo "== batch $1 path $2"; DISB200_TV_PATH=$2 timeout 300 python bench.py --steps 10 --warmup 3 --batch $1 --no-e2e --no-cpu-baseline 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print(round(d['value']), round(d['ms_per_step'],3), d['gpu_launches'], d['roofline']['stage_ms_per_step'])
    else: print(l.rstrip())"
done
