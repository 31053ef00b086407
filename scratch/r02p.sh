#!/bin/bash
# round 2, call p (1 GPU): what decides staged against direct tap loads in the 8x8 grey search - overlap (grid step) or iteration count?
O=gpurun_out
run() {
  timeout 600 python bench.py "$@" --warmup 3 --steps 10 --batch 1776 --no-e2e --no-cpu-baseline --no-parity 2>/dev/null | grep '^{' | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l); print(round(d['value'],1), 'pairs/s', 'search ms', d['roofline']['stage_ms_per_step']['inverse_search'], d['config'].get('param_overrides'))"
}
: > $O/r02p_stage_params.txt
for cfg in "--workload op1 --set patove=0.4" "--workload op1 --set max_iter=12,min_iter=12" "--workload C1 --set patove=0.3" "--workload C1 --set max_iter=16,min_iter=16" "--workload C1 --set usetvref=0" "--workload op1 --set usetvref=1" "--workload C1 --set patove=0.5" "--workload C1 --set patove=0.65"; do
  for pm in 0 1; do
    echo -n "[patch_map=$pm] $cfg: " | tee -a $O/r02p_stage_params.txt
    run --patch-map $pm $cfg | tee -a $O/r02p_stage_params.txt
  done
done
