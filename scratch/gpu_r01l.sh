#!/bin/bash
mkdir -p gpurun_out
echo "== new tests"; timeout 600 python -m pytest tests -m gpu -x -q -k "gradmag or lowres or cli_matches" 2>&1 | tail -6
echo "== full suite"; timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -4
echo "== smoke"; timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -3
echo "== bench default"; timeout 600 python bench.py --steps 20 --warmup 3 > gpurun_out/bench_r01l.json 2> gpurun_out/bench_r01l.err; tail -3 gpurun_out/bench_r01l.err
python - <<'PY'
import json
d=json.loads([l for l in open("gpurun_out/bench_r01l.json") if l.startswith("{")][0])
print(d["value"], d["ms_per_step"], d["e2e"], d["e2e_ofclass_cut"], d["gpu_launches"], d["cpu_baseline"])
print(d["roofline"]["stage_ms_per_step"]); print(d["parity"]); print(d["clocks"])
PY
for c in 64 111 444; do echo "== lowres e2e chunk $c"; DISB200_HOST_CHUNK=$c timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print(d['e2e']['value'], d['e2e_ofclass_cut']['value'])"; done
echo "== reference arm"; timeout 300 python bench.py --impl reference --steps 3 --warmup 1 2>&1 | tail -1 | cut -c1-600
