#!/bin/bash
# round 2, call af (1 GPU): 8x8 BGR search with staged tap windows copied only when they moved, against loads from global memory
# (libdisb200_nobgr.so)
O=gpurun_out
OLD=$PWD/of_dis_b200/libdisb200_nobgr.so
echo "== parity"; timeout 150 python -m pytest tests/test_parity_gpu.py -m gpu -q -x -k "of_rgb_op2 or de_rgb_op2 or of_rgb_l1_fb or ofclass_against_reference_golden or compiled_reference" 2>&1 | tail -3
run() {
  timeout 100 python bench.py "$@" --workload rgb2 --batch 1184 --warmup 3 --steps 8 --no-e2e --no-cpu-baseline 2>/dev/null | grep '^{' | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l); p=d.get('parity') or {}; print(round(d['value'],1), 'pairs/s', round(d['ms_per_step'],3), 'ms', 'search ms', d['roofline']['stage_ms_per_step']['inverse_search'], 'EPE gpu-cpu mean/max', p.get('epe_gpu_vs_cpu_mean_px'), p.get('epe_gpu_vs_cpu_max_px'))"
}
: > $O/r02af_stage_bgr.txt
echo -n "[staged] rgb2: " | tee -a $O/r02af_stage_bgr.txt; run | tee -a $O/r02af_stage_bgr.txt
echo -n "[global] rgb2: " | tee -a $O/r02af_stage_bgr.txt; DISB200_LIB=$OLD run --no-parity | tee -a $O/r02af_stage_bgr.txt
