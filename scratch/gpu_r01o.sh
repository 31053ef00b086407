#!/bin/bash
mkdir -p gpurun_out
run() { # workload batch steps
  timeout 400 python bench.py --workload $1 --batch $2 --steps $3 --warmup 3 --no-e2e --cpu-seconds 4 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); p=d['parity']; print(round(d['value'],1), 'pairs/s', round(d['ms_per_step'],3), 'ms', 'search', d['roofline']['stage_ms_per_step']['inverse_search'], 'cpu', round(d['cpu_baseline']['value'],2), 'epe mean %.3g max %.3g gt %.6f / %.6f' % (p['epe_gpu_vs_cpu_mean_px'], p['epe_gpu_vs_cpu_max_px'], p['epe_vs_ground_truth_gpu_px'], p['epe_vs_ground_truth_cpu_px']))
    else: print(l.rstrip()[-200:])"
}
echo "== rgb2 value-interleaved"; run rgb2 1184 6
echo "== rgb2 block"; DISB200_PATCH_RGB8=b run rgb2 1184 6
echo "== rgb2 block+fma"; DISB200_PATCH_RGB8=bf run rgb2 1184 6
echo "== tests rgb block"; DISB200_PATCH_RGB8=b timeout 600 python -m pytest tests -m gpu -x -q -k "rgb" 2>&1 | tail -3
echo "== tests rgb block+fma"; DISB200_PATCH_RGB8=bf timeout 600 python -m pytest tests -m gpu -x -q -k "rgb" 2>&1 | tail -3
echo "== defaults: tests psz12"; timeout 600 python -m pytest tests -m gpu -x -q -k "op3 or op4 or psz12 or c2 or golden" 2>&1 | tail -3
echo "== op4 default"; run op4 148 3
echo "== C2 default"; run C2 64 2
