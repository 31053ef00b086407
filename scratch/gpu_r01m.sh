#!/bin/bash
mkdir -p gpurun_out
echo "== tests (block map)"; timeout 900 python -m pytest tests -m gpu -x -q -k "op3 or op4 or psz12 or c2 or golden or run_dense_parity or lowres" 2>&1 | tail -5
run() { # workload batch steps
  timeout 400 python bench.py --workload $1 --batch $2 --steps $3 --warmup 3 --no-e2e --cpu-seconds 8 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print(round(d['value'],1), 'pairs/s', round(d['ms_per_step'],3), 'ms', d['config']['pairs_per_step_per_gpu'], 'cpu', d['cpu_baseline'] and round(d['cpu_baseline']['value'],2), d['parity'], d['roofline']['stage_ms_per_step'])
    else: print(l.rstrip()[-200:])"
}
for w in "op3 592 6" "op4 148 4" "C2 64 3"; do
  set -- $w
  echo "== $1 block map"; run $1 $2 $3
  echo "== $1 value-interleaved (old)"; DISB200_PATCH_MAP=v run $1 $2 $3
done
