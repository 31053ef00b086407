#!/bin/bash
# round 2, call c: full GPU suite incl. tiled (virtual strips) + warm start fix
mkdir -p gpurun_out
echo "== tiled tests first"; timeout 900 python -m pytest tests/test_parity_gpu.py -m gpu -x -q -k "tiled or warm_start" 2>&1 | tail -15
echo "== full suite"; timeout 2400 python -m pytest tests -m gpu -q 2>&1 | tail -15
