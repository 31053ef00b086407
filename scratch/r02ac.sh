#!/bin/bash
# round 2, call ac (1 GPU): the bench line after the median-of-5 stage timing change
timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline 2>&1 | tail -1 | python -c "
import sys,json
d=json.loads(sys.stdin.read()); print(d['value'], d['ms_per_step'], d['e2e']['value'], d['roofline']['kernel'], d['roofline']['frac'], d['roofline']['stage_ms_per_step'], d['parity']['bit_identical'])"
