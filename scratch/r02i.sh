#!/bin/bash
# round 2, call i (1 GPU): 12-pixel SOR ring (8 pairs per SM) against the 16-pixel ring (6 per SM): parity, then A/B timings
O=gpurun_out
echo "== tv paths + whole path"; timeout 900 python -m pytest tests/test_parity_gpu.py -m gpu -q -x -k "tv_paths or C1_of_int_op2 or C3_de or of_rgb_op2 or full_batch or sweep_counts or varref_level_bit" 2>&1 | tail -4
run() {
  timeout 300 python bench.py "$@" --warmup 3 --no-e2e --no-cpu-baseline 2>/dev/null | grep '^{' | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l); p=d.get('parity'); print(round(d['value'],1), 'pairs/s', round(d['ms_per_step'],3), 'ms', d['config']['pairs_per_step_per_gpu'], d['gpu_launches'], d['roofline']['stage_ms_per_step'], p and p['bit_identical'])"
}
: > $O/r02i_ring.txt
for cfg in "--batch 1776" "--batch 2368" "--batch 1184" "--streams 1 --batch 1184" "--streams 1 --batch 2368" "--batch 4736" "--workload C3 --batch 2368" "--workload C3 --batch 1776" "--workload C5 --batch 592" "--workload rgb2 --batch 1184"; do
  echo -n "[ring12] $cfg: " | tee -a $O/r02i_ring.txt
  run --steps 10 $cfg | tee -a $O/r02i_ring.txt
done
for cfg in "--batch 1776" "--batch 2368" "--workload C3 --batch 1776" "--workload C5 --batch 592" "--workload rgb2 --batch 1184"; do
  echo -n "[ring16] $cfg: " | tee -a $O/r02i_ring.txt
  DISB200_LIB=$PWD/of_dis_b200/libdisb200_ring16.so run --steps 10 $cfg | tee -a $O/r02i_ring.txt
done
