#!/bin/bash
echo "== bench default"; timeout 600 python bench.py --steps 20 --warmup 3 > gpurun_out/bench_r01k.json 2> gpurun_out/bench_r01k.err; tail -2 gpurun_out/bench_r01k.err
python - <<'PY'
import json
d=json.loads([l for l in open("gpurun_out/bench_r01k.json") if l.startswith("{")][0])
r=d["roofline"]; print(d["value"], d["ms_per_step"], d["e2e"]["value"], d["gpu_launches"], d["config"]["pairs_per_step_per_gpu"], d["cpu_baseline"]["value"])
print({k:r[k] for k in ("kernel","stage","achieved","frac","traffic","share_of_step")})
print(r["stage_ms_per_step"]); print(d["parity"])
PY
for w in C5 C3 C4 op1; do echo "== $w"; timeout 300 python bench.py --workload $w --steps 6 --warmup 3 --no-e2e --cpu-seconds 8 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print(round(d['value'],1), round(d['ms_per_step'],3), d['config']['pairs_per_step_per_gpu'], 'cpu', round(d['cpu_baseline']['value'],2), d['cpu_baseline']['sample'][:60], d['parity'])
    else: print(l.rstrip()[-300:])"; done
DISB200_STREAMS=1 timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -s 204 -c 136 --csv --log-file gpurun_out/launches_r01k.csv python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu-baseline > gpurun_out/ncu_launch_r01k.log 2>&1; echo "launch list rc=$?"
