#!/bin/bash
# round 2, call ae (1 GPU): sanity of the 8x8 paths and smoke after the staged 12x12 change
timeout 200 python -m pytest tests/test_parity_gpu.py tests/test_zz_no_alloc.py -m gpu -q -x -k "C1_of_int_op2 or staged or full_batch or op1 or C3_de or no_alloc or static" 2>&1 | tail -3
timeout 100 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
