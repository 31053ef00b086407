#!/bin/bash
timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -4
echo "== streams=2 suite"; DISB200_STREAMS=2 timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
echo "== bench default"; timeout 600 python bench.py --steps 20 --warmup 3 > gpurun_out/bench_r01h.json 2> gpurun_out/bench_r01h.err; tail -c 1500 gpurun_out/bench_r01h.json; tail -3 gpurun_out/bench_r01h.err
echo "== smoke"; timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -3
