#!/bin/bash
# round 2, call z (1 GPU): staged search that copies a window again only when its integer origin moved, against copying every
# evaluation (libdisb200_noskip.so); the new tv_paths cases
O=gpurun_out
NS=$PWD/of_dis_b200/libdisb200_noskip.so
echo "== parity"; timeout 900 python -m pytest tests/test_parity_gpu.py -m gpu -q -x -k "run_dense_parity or compiled_reference or staged or full_batch or tv_paths" 2>&1 | tail -4
run() {
  timeout 600 python bench.py "$@" --warmup 3 --steps 10 --no-e2e --no-cpu-baseline 2>/dev/null | grep '^{' | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l); p=d.get('parity') or {}; print(round(d['value'],1), 'pairs/s', round(d['ms_per_step'],3), 'ms', d['config']['pairs_per_step_per_gpu'], 'search ms', d['roofline']['stage_ms_per_step']['inverse_search'], 'bit-identical', p.get('bit_identical'))"
}
: > $O/r02z_stage_skip.txt
for rep in 1 2; do
for cfg in "--batch 1776" "--workload C3 --batch 1776" "--workload C5 --batch 592"; do
  echo -n "[skip unchanged] $cfg: " | tee -a $O/r02z_stage_skip.txt
  run $cfg | tee -a $O/r02z_stage_skip.txt
  echo -n "[copy always] $cfg: " | tee -a $O/r02z_stage_skip.txt
  DISB200_LIB=$NS run $cfg | tee -a $O/r02z_stage_skip.txt
done
done
