#!/bin/bash
# round 2, call l (1 GPU): 8x8 grey search with staged tap windows (patch_map=2): parity, then A/B timings
O=gpurun_out
echo "== staged parity"; timeout 600 python -m pytest tests/test_parity_gpu.py -m gpu -q -x -k "staged_tap" 2>&1 | tail -4
run() {
  timeout 300 python bench.py "$@" --warmup 3 --no-e2e --no-cpu-baseline 2>/dev/null | grep '^{' | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l); p=d.get('parity'); print(round(d['value'],1), 'pairs/s', round(d['ms_per_step'],3), 'ms', d['config']['pairs_per_step_per_gpu'], d['gpu_launches'], d['roofline']['stage_ms_per_step'], p and p['bit_identical'])"
}
: > $O/r02l_stage.txt
for pm in 2 0 2 0; do
for cfg in "--batch 1776" "--workload op1 --batch 1776" "--workload C3 --batch 1776" "--workload C5 --batch 592"; do
  echo -n "[patch_map=$pm] $cfg: " | tee -a $O/r02l_stage.txt
  run --steps 10 --patch-map $pm $cfg | tee -a $O/r02l_stage.txt
done
done
