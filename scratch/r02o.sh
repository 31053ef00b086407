#!/bin/bash
# round 2, call o (1 GPU): (1) staged against direct tap loads again, now that the direct-load kernels are back at 128 registers
# (launch bounds without an occupancy request); (2) reciprocal products in the solve of the tolerance-compared search kernels
# against exact divisions (libdisb200_exactdiv.so); (3) C5 at 256 pairs, fused against split refinement
O=gpurun_out
ED=$PWD/of_dis_b200/libdisb200_exactdiv.so
echo "== parity"; timeout 900 python -m pytest tests/test_parity_gpu.py -m gpu -q -x -k "run_dense_parity or compiled_reference or c4_heavy or staged" 2>&1 | tail -4
run() {
  timeout 600 python bench.py "$@" --warmup 3 --no-e2e --no-cpu-baseline --no-parity 2>/dev/null | grep '^{' | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l); print(round(d['value'],1), 'pairs/s', round(d['ms_per_step'],3), 'ms', d['config']['pairs_per_step_per_gpu'], d['roofline']['stage_ms_per_step'])"
}
: > $O/r02o_stage.txt
for pm in 0 1 0 1; do
for cfg in "--batch 1776" "--workload op1 --batch 1776" "--workload C3 --batch 1776" "--workload C5 --batch 592"; do
  echo -n "[patch_map=$pm] $cfg: " | tee -a $O/r02o_stage.txt
  run --steps 10 --patch-map $pm $cfg | tee -a $O/r02o_stage.txt
done
done
: > $O/r02o_rcp.txt
for cfg in "--workload op4 --batch 148 --steps 4" "--workload op3 --batch 592 --steps 6" "--workload C2 --batch 64 --steps 3"; do
  echo -n "[rcp] $cfg: " | tee -a $O/r02o_rcp.txt
  run $cfg | tee -a $O/r02o_rcp.txt
  echo -n "[div] $cfg: " | tee -a $O/r02o_rcp.txt
  DISB200_LIB=$ED run $cfg | tee -a $O/r02o_rcp.txt
done
: > $O/r02o_c5_b256.txt
for ts in 1 2 1 2; do
  echo -n "[tv_schedule=$ts] C5 256: " | tee -a $O/r02o_c5_b256.txt
  run --workload C5 --batch 256 --steps 10 --tv-schedule $ts | tee -a $O/r02o_c5_b256.txt
done
