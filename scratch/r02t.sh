#!/bin/bash
# round 2, call t (1 GPU): team SOR (one warp per band, one CTA per pair, lock-step through a shared-memory mailbox) against the
# one-warp-per-pair SOR: bit-identity (tv_paths), then A/B timings
O=gpurun_out
echo "== tv paths"; timeout 900 python -m pytest tests/test_parity_gpu.py -m gpu -q -x -k "tv_paths" 2>&1 | tail -4
run() {
  timeout 600 python bench.py "$@" --warmup 3 --steps 10 --no-e2e --no-cpu-baseline 2>/dev/null | grep '^{' | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l); p=d.get('parity') or {}; print(round(d['value'],1), 'pairs/s', round(d['ms_per_step'],3), 'ms', d['config']['pairs_per_step_per_gpu'], 'sor ms', d['roofline']['stage_ms_per_step']['sor'], 'bit-identical', p.get('bit_identical'))"
}
: > $O/r02t_team.txt
for cfg in "--batch 1776" "--workload C3 --batch 1776" "--workload C5 --batch 592" "--workload rgb2 --batch 1184" "--batch 888 --streams 1" "--batch 296 --streams 1" "--workload C5 --batch 256"; do
  for ts in 5 2; do
    echo -n "[tv_schedule=$ts] $cfg: " | tee -a $O/r02t_team.txt
    run --tv-schedule $ts $cfg | tee -a $O/r02t_team.txt
  done
done
