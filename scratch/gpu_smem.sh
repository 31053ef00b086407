#!/bin/bash
timeout 900 python -m pytest tests/test_parity_gpu.py -m gpu -x -q -k "tv_paths or varref or full_batch or run_dense" 2>&1 | tail -3
SKIPTEST=1 bash scratch/gpu_quick.sh "1480 split" "1776 split" "2368 split"
