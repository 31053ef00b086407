#!/bin/bash
mkdir -p gpurun_out
run() { # workload batch steps
  timeout 300 python bench.py --workload $1 --batch $2 --steps $3 --warmup 3 --no-e2e --no-cpu-baseline 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print(round(d['value'],1), 'pairs/s', round(d['ms_per_step'],3), 'ms', d['gpu_launches'], d['roofline']['stage_ms_per_step'])
    else: print(l.rstrip()[-200:])"
}
for b in 16 32 64 96; do echo "== C4 batch $b"; run C4 $b 2; done
echo "== C4 batch 64, band mode forced"; DISB200_SORB_MAX=100000 run C4 64 2
nvidia-smi --query-gpu=memory.used,memory.total --format=csv
free -g | head -2
