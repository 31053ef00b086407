#!/bin/bash
timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
timeout 600 python bench.py --steps 10 --warmup 3 --no-e2e > gpurun_out/bench_ps.json 2> gpurun_out/bench_ps.err; tail -2 gpurun_out/bench_ps.err
python - <<'PY'
import json
d=json.loads([l for l in open("gpurun_out/bench_ps.json") if l.startswith("{")][0])
r=d["roofline"]; print(d["value"], d["ms_per_step"], d["gpu_launches"], d["config"]["pairs_per_step_per_gpu"])
print(r["stage_ms_per_step"]); print(d["parity"])
PY
