#!/bin/bash
# round 2, call ad (1 GPU): packed 12x12 grey search with staged tap windows copied only when they moved, against loads from global
# memory (libdisb200_nostage12.so)
O=gpurun_out
OLD=$PWD/of_dis_b200/libdisb200_nostage12.so
echo "== parity"; timeout 300 python -m pytest tests/test_parity_gpu.py -m gpu -q -x -k "of_int_op3 or de_int_fb_op3 or compiled_reference or c4_heavy" 2>&1 | tail -3
run() {
  timeout 200 python bench.py "$@" --warmup 2 --no-e2e --no-cpu-baseline 2>/dev/null | grep '^{' | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l); p=d.get('parity') or {}; print(round(d['value'],1), 'pairs/s', round(d['ms_per_step'],3), 'ms', d['config']['pairs_per_step_per_gpu'], 'search ms', d['roofline']['stage_ms_per_step']['inverse_search'], 'EPE gpu-cpu mean/max', p.get('epe_gpu_vs_cpu_mean_px'), p.get('epe_gpu_vs_cpu_max_px'))"
}
: > $O/r02ad_stage12.txt
echo -n "[staged] op4: " | tee -a $O/r02ad_stage12.txt; run --workload op4 --batch 148 --steps 3 | tee -a $O/r02ad_stage12.txt
echo -n "[global] op4: " | tee -a $O/r02ad_stage12.txt; DISB200_LIB=$OLD run --workload op4 --batch 148 --steps 3 --no-parity | tee -a $O/r02ad_stage12.txt
echo -n "[staged] op3: " | tee -a $O/r02ad_stage12.txt; run --workload op3 --batch 592 --steps 5 --no-parity | tee -a $O/r02ad_stage12.txt
echo -n "[global] op3: " | tee -a $O/r02ad_stage12.txt; DISB200_LIB=$OLD run --workload op3 --batch 592 --steps 5 --no-parity | tee -a $O/r02ad_stage12.txt
