#!/bin/bash
mkdir -p gpurun_out
run() { # workload batch steps
  timeout 400 python bench.py --workload $1 --batch $2 --steps $3 --warmup 3 --no-e2e --cpu-seconds 4 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); p=d['parity']; print(round(d['value'],1), 'pairs/s', round(d['ms_per_step'],3), 'ms', 'search', d['roofline']['stage_ms_per_step']['inverse_search'], 'epe mean %.3g max %.3g gt %.6f / %.6f' % (p['epe_gpu_vs_cpu_mean_px'], p['epe_gpu_vs_cpu_max_px'], p['epe_vs_ground_truth_gpu_px'], p['epe_vs_ground_truth_cpu_px']))
    else: print(l.rstrip()[-200:])"
}
echo "== op4 block"; run op4 148 3
echo "== op4 block+fma"; DISB200_PATCH_FMA=1 run op4 148 3
echo "== op4 g4"; DISB200_PATCH_G4=1 run op4 148 3
echo "== op4 g4+fma"; DISB200_PATCH_G4=1 DISB200_PATCH_FMA=1 run op4 148 3
echo "== op3 block+fma"; DISB200_PATCH_FMA=1 run op3 592 4
echo "== op3 g4+fma"; DISB200_PATCH_G4=1 DISB200_PATCH_FMA=1 run op3 592 4
echo "== C2 block+fma"; DISB200_PATCH_FMA=1 run C2 64 2
echo "== tests with fma"; DISB200_PATCH_FMA=1 timeout 600 python -m pytest tests -m gpu -x -q -k "op3 or op4 or psz12 or c2 or golden" 2>&1 | tail -3
echo "== ncu op4 level-0 search (block map, no fma)"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:patch_search --launch-skip 5 -c 1 -o gpurun_out/r01n_search12 -f python bench.py --workload op4 --batch 16 --steps 1 --warmup 3 --no-e2e --no-cpu-baseline > gpurun_out/ncu_r01n.log 2>&1; echo "ncu rc=$?"; tail -2 gpurun_out/ncu_r01n.log
