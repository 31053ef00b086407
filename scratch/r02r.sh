#!/bin/bash
# round 2, call r (1 GPU): 12x12 grey search with packed fp32 pairs (FFMA2 / FADD2 / FMUL2) against the scalar FM kernel
# (libdisb200_nopk.so)
O=gpurun_out
NP=$PWD/of_dis_b200/libdisb200_nopk.so
echo "== parity"; timeout 900 python -m pytest tests/test_parity_gpu.py -m gpu -q -x -k "run_dense_parity or compiled_reference or c4_heavy" 2>&1 | tail -4
run() {
  timeout 600 python bench.py "$@" --warmup 3 --no-e2e --no-cpu-baseline 2>/dev/null | grep '^{' | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l); p=d.get('parity') or {}; print(round(d['value'],1), 'pairs/s', round(d['ms_per_step'],3), 'ms', d['config']['pairs_per_step_per_gpu'], 'search ms', d['roofline']['stage_ms_per_step']['inverse_search'], 'EPE gpu-cpu mean/max', p.get('epe_gpu_vs_cpu_mean_px'), p.get('epe_gpu_vs_cpu_max_px'))"
}
: > $O/r02r_packed.txt
for cfg in "--workload op4 --batch 148 --steps 4" "--workload op3 --batch 592 --steps 6"; do
  echo -n "[packed] $cfg: " | tee -a $O/r02r_packed.txt
  run $cfg | tee -a $O/r02r_packed.txt
  echo -n "[scalar] $cfg: " | tee -a $O/r02r_packed.txt
  DISB200_LIB=$NP run $cfg | tee -a $O/r02r_packed.txt
done
