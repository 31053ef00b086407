#!/bin/bash
# round 2, call n (1 GPU): repeat detection in the inverse search (skip forced iterations that replay a fixed point / 2-cycle):
# bit-identity against a build without it, parity tests, A/B timings on op4 / op3 / C2 / rgb2 / C4
O=gpurun_out
NO=$PWD/of_dis_b200/libdisb200_nocyc.so
python scratch/ab_identity.py /tmp/ab_new.npz && DISB200_LIB=$NO python scratch/ab_identity.py /tmp/ab_old.npz && python scratch/ab_identity.py --cmp /tmp/ab_new.npz /tmp/ab_old.npz 2>&1 | tee $O/r02n_identity.txt
echo "== parity"; timeout 900 python -m pytest tests/test_parity_gpu.py -m gpu -q -x -k "run_dense_parity or compiled_reference or c4_heavy" 2>&1 | tail -4
run() {
  timeout 600 python bench.py "$@" --warmup 2 --no-e2e --no-cpu-baseline --no-parity 2>/dev/null | grep '^{' | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l); print(round(d['value'],1), 'pairs/s', round(d['ms_per_step'],3), 'ms', d['config']['pairs_per_step_per_gpu'], 'search ms', d['roofline']['stage_ms_per_step']['inverse_search'])"
}
: > $O/r02n_repeat_skip.txt
for cfg in "--workload op4 --batch 148 --steps 4" "--workload op3 --batch 592 --steps 6" "--workload C2 --batch 64 --steps 3" "--workload rgb2 --batch 1184 --steps 8" "--workload C4 --batch 148 --steps 3"; do
  echo -n "[skip] $cfg: " | tee -a $O/r02n_repeat_skip.txt
  run $cfg | tee -a $O/r02n_repeat_skip.txt
  echo -n "[none] $cfg: " | tee -a $O/r02n_repeat_skip.txt
  DISB200_LIB=$NO run $cfg | tee -a $O/r02n_repeat_skip.txt
done
