#!/bin/bash
# round 2, call ag (1 GPU): 12x12 BGR search staged (244 registers, libdisb200_bgr12.so) against global loads (the default), C2
O=gpurun_out
NEW=$PWD/of_dis_b200/libdisb200_bgr12.so
DISB200_LIB=$NEW timeout 60 python -m pytest tests/test_parity_gpu.py -m gpu -q -x -k "of_rgb_op3" 2>&1 | tail -2
run() {
  timeout 60 python bench.py --workload C2 --batch 64 --warmup 1 --steps 2 --no-e2e --no-cpu-baseline --no-parity 2>/dev/null | grep '^{' | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l); print(round(d['value'],1), 'pairs/s', round(d['ms_per_step'],3), 'ms', 'search ms', d['roofline']['stage_ms_per_step']['inverse_search'])"
}
: > $O/r02ag_stage_bgr12.txt
echo -n "[staged, 244 registers] C2: " | tee -a $O/r02ag_stage_bgr12.txt; DISB200_LIB=$NEW run | tee -a $O/r02ag_stage_bgr12.txt
echo -n "[global loads] C2: " | tee -a $O/r02ag_stage_bgr12.txt; run | tee -a $O/r02ag_stage_bgr12.txt
