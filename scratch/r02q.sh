#!/bin/bash
# round 2, call q (1 GPU): FM search kernels with reciprocal means + squared outlier test on top of the reciprocal solve
# (libdisb200_rcp1.so = solve only); staged / direct selection by grid step
O=gpurun_out
R1=$PWD/of_dis_b200/libdisb200_rcp1.so
echo "== parity"; timeout 900 python -m pytest tests/test_parity_gpu.py -m gpu -q -x -k "run_dense_parity or compiled_reference or c4_heavy or staged" 2>&1 | tail -4
run() {
  timeout 600 python bench.py "$@" --warmup 3 --no-e2e --no-cpu-baseline --no-parity 2>/dev/null | grep '^{' | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l); print(round(d['value'],1), 'pairs/s', round(d['ms_per_step'],3), 'ms', d['config']['pairs_per_step_per_gpu'], 'search ms', d['roofline']['stage_ms_per_step']['inverse_search'])"
}
: > $O/r02q_rcp2.txt
for cfg in "--workload op4 --batch 148 --steps 4" "--workload op3 --batch 592 --steps 6" "--workload C2 --batch 64 --steps 3"; do
  echo -n "[rcp solve+mean+outlier] $cfg: " | tee -a $O/r02q_rcp2.txt
  run $cfg | tee -a $O/r02q_rcp2.txt
  echo -n "[rcp solve only] $cfg: " | tee -a $O/r02q_rcp2.txt
  DISB200_LIB=$R1 run $cfg | tee -a $O/r02q_rcp2.txt
done
for cfg in "--workload op1 --batch 1776 --steps 10" "--batch 1776 --steps 10"; do
  echo -n "[auto] $cfg: " | tee -a $O/r02q_rcp2.txt
  run $cfg | tee -a $O/r02q_rcp2.txt
done
