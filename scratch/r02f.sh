#!/bin/bash
# round 2, call f (1 GPU): system assembly inside the one-warp SOR kernel - parity of all schedules, then A/B timings
O=gpurun_out
echo "== tv paths"; timeout 900 python -m pytest tests/test_parity_gpu.py -m gpu -q -x -k "tv_paths or C1_of_int_op2 or C3_de or of_rgb_op2 or full_batch" 2>&1 | tail -6
run() {
  timeout 300 python bench.py "$@" --warmup 3 --no-e2e --no-cpu-baseline 2>/dev/null | grep '^{' | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l); p=d.get('parity'); print(round(d['value'],1), 'pairs/s', round(d['ms_per_step'],3), 'ms', d['config']['pairs_per_step_per_gpu'], d['gpu_launches'], d['roofline']['stage_ms_per_step'], p and p['bit_identical'])"
}
: > $O/r02f_solvef.txt
for cfg in "--tv-schedule 2" "" "--batch 1480" "--batch 1480 --tv-schedule 2" "--streams 1 --batch 740" "--streams 1 --batch 1480" "--batch 2960" "--batch 2960 --tv-schedule 2" "--workload C3" "--workload C3 --tv-schedule 2" "--workload rgb2 --batch 1184" "--workload rgb2 --batch 1184 --tv-schedule 2" "--workload C5 --batch 592" "--workload C5 --batch 592 --tv-schedule 2"; do
  echo -n "[128 regs] $cfg: " | tee -a $O/r02f_solvef.txt
  run --steps 10 $cfg | tee -a $O/r02f_solvef.txt
done
for cfg in "" "--batch 1480" "--batch 2664"; do
  echo -n "[112 regs] $cfg: " | tee -a $O/r02f_solvef.txt
  DISB200_LIB=$PWD/of_dis_b200/libdisb200_r112.so run --steps 10 $cfg | tee -a $O/r02f_solvef.txt
done
