#!/bin/bash
timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
echo "== bench default"; timeout 600 python bench.py --steps 20 --warmup 3 > gpurun_out/bench_r01j.json 2> gpurun_out/bench_r01j.err; tail -2 gpurun_out/bench_r01j.err
python - <<'PY'
import json
d=json.loads([l for l in open("gpurun_out/bench_r01j.json") if l.startswith("{")][0])
r=d["roofline"]; print(d["value"], d["ms_per_step"], d["e2e"]["value"], d["gpu_launches"], d["config"]["pairs_per_step_per_gpu"], d["cpu_baseline"]["value"])
print({k:r[k] for k in ("kernel","stage","achieved","frac","traffic","share_of_step")})
print(r["stage_ms_per_step"]); print(d["parity"])
PY
echo "== reference arm"; timeout 300 python bench.py --impl reference --steps 3 --warmup 1 | cut -c1-400
echo "== smoke"; timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
