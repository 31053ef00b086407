#!/bin/bash
# round 2, call d (1 GPU): failing tests + anti-phase experiment
mkdir -p gpurun_out
O=gpurun_out
echo "== tests"; timeout 900 python -m pytest tests/test_parity_gpu.py -m gpu -q -k "antiphase or warm_start or initflow or tiled_small" 2>&1 | tail -8
run() {
  timeout 300 python bench.py "$@" --warmup 3 --no-e2e --no-cpu-baseline --no-parity 2>/dev/null | grep '^{' | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l); print(round(d['value'],1), 'pairs/s', round(d['ms_per_step'],3), 'ms', d['config']['pairs_per_step_per_gpu'], d['gpu_launches'])"
}
: > $O/r02d_antiphase.txt
for cfg in "--streams 2" "--streams 2 --antiphase 1" "--streams 1" "--streams 2 --antiphase 1 --batch 3552" "--streams 2 --batch 3552" "--streams 2 --antiphase 1 --batch 1184" "--streams 2 --batch 1184" "--streams 2 --antiphase 1 --workload C5 --batch 592" "--streams 2 --workload C5 --batch 592" "--streams 2 --antiphase 1 --workload rgb2 --batch 1184" "--workload rgb2 --batch 1184" "--streams 2 --antiphase 1 --workload C3" "--workload C3"; do
  echo -n "$cfg: " | tee -a $O/r02d_antiphase.txt
  run --steps 10 $cfg | tee -a $O/r02d_antiphase.txt
done
