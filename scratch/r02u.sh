#!/bin/bash
# round 2, call u (1 GPU): where does the fused refinement (one CTA per pair runs a whole level) still beat the split path now that
# the split path's SOR is the team schedule?  batch sweep, fused (1) against split + team (5), single stream
O=gpurun_out
run() {
  timeout 300 python bench.py "$@" --warmup 3 --steps 20 --streams 1 --no-e2e --no-cpu-baseline --no-parity 2>/dev/null | grep '^{' | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l); print(round(d['value'],1), 'pairs/s', round(d['ms_per_step'],3), 'ms', d['config']['pairs_per_step_per_gpu'], 'sor+system+smooth ms', round(d['roofline']['stage_ms_per_step']['sor']+d['roofline']['stage_ms_per_step']['system']+d['roofline']['stage_ms_per_step']['smooth'],4))"
}
: > $O/r02u_fused_vs_team.txt
for wl in C1 C5 C3; do
for b in 1 8 32 64 148 222 296; do
  for ts in 1 5; do
    echo -n "[$wl tv_schedule=$ts] batch $b: " | tee -a $O/r02u_fused_vs_team.txt
    run --workload $wl --tv-schedule $ts --batch $b | tee -a $O/r02u_fused_vs_team.txt
  done
done
done
