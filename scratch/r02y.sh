#!/bin/bash
# round 2, call y (2 GPUs): the physical multi-GPU tiling tests with the final build
timeout 900 python -m pytest tests/test_tiled_multi_gpu.py tests/test_zz_no_alloc.py -m gpu -q 2>&1 | tail -5
