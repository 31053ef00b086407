#!/bin/bash
# round 2, call e (N GPUs): one pair tiled over physical GPUs - bit-identity tests and the 4K latency table
# usage: bash scratch/r02e.sh <ngpu>
N=${1:-2}
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/r02e_topo_$N.txt 2>&1
echo "== multi-GPU tiled tests ($N GPUs)"; timeout 1200 python -m pytest tests/test_tiled_multi_gpu.py -m gpu -q 2>&1 | tail -8
echo "== 4K latency table"; timeout 1200 python tools/tiled_latency.py --out gpurun_out/r02e_tiled_latency_$N.json --reps 5 $( [ "$N" = 2 ] && echo --virtual 2 ) 2>&1 | grep -v "^{" | tail -20
