#!/bin/bash
mkdir -p gpurun_out
echo "== full suite"; timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
echo "== smoke"; timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
echo "== bench default"; timeout 600 python bench.py --steps 20 --warmup 3 > gpurun_out/bench_r01q.json 2> gpurun_out/bench_r01q.err; tail -2 gpurun_out/bench_r01q.err
python - <<'PY'
import json
d=json.loads([l for l in open("gpurun_out/bench_r01q.json") if l.startswith("{")][0])
r=d["roofline"]; print(d["value"], d["ms_per_step"], d["e2e"]["value"], d["e2e_ofclass_cut"]["value"], d["gpu_launches"], d["cpu_baseline"]["value"])
print({k:r[k] for k in ("kernel","stage","achieved","frac","traffic","share_of_step")}); print(r["whole_path"])
print(r["stage_ms_per_step"]); print(d["parity"]); print(d["clocks"])
PY
: > gpurun_out/workloads_r01q.jsonl
for w in "rgb2 1184 6" "op1 1776 10" "C3 1776 10" "C5 256 6" "op3 592 4" "op4 148 3" "C2 64 2" "C4 148 2"; do set -- $w; echo "== $1"; timeout 400 python bench.py --workload $1 --batch $2 --steps $3 --warmup 3 --no-e2e --cpu-seconds 6 2>/dev/null | grep '^{' | tee -a gpurun_out/workloads_r01q.jsonl | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l); p=d['parity']; print(round(d['value'],1), 'pairs/s', round(d['ms_per_step'],3), 'ms', d['config']['pairs_per_step_per_gpu'], 'cpu', round(d['cpu_baseline']['value'],2), 'bit' if p['bit_identical'] else 'epe mean %.3g max %.3g' % (p['epe_gpu_vs_cpu_mean_px'], p['epe_gpu_vs_cpu_max_px']), d['roofline']['stage_ms_per_step'])"; done
echo "== launch list"; DISB200_STREAMS=1 timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -s 204 -c 136 --csv --log-file gpurun_out/launches_r01q.csv python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu-baseline > gpurun_out/ncu_launch_r01q.log 2>&1; echo "launch list rc=$?"
echo "== reference arm"; timeout 300 python bench.py --impl reference --steps 3 --warmup 1 2>&1 | tail -1 | cut -c1-300
