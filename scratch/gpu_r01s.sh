#!/bin/bash
mkdir -p gpurun_out
echo "== full suite"; timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
echo "== smoke"; timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
echo "== bench default"; timeout 600 python bench.py --steps 20 --warmup 3 > gpurun_out/bench_r01s.json 2> gpurun_out/bench_r01s.err; tail -2 gpurun_out/bench_r01s.err
python - <<'PY'
import json
d=json.loads([l for l in open("gpurun_out/bench_r01s.json") if l.startswith("{")][0])
print(d["value"], d["ms_per_step"], d["e2e"]["value"], d["e2e_ofclass_cut"]["value"], d["gpu_launches"], d["cpu_baseline"]["value"], d["parity"]["bit_identical"], d["clocks"])
PY
run() { # args...
  timeout 400 python bench.py "$@" --warmup 3 --no-e2e 2>/dev/null | grep '^{' | tee -a gpurun_out/extra_r01s.jsonl | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l); p=d.get('parity'); print(round(d['value'],1), 'pairs/s', round(d['ms_per_step'],3), 'ms', d['config']['pairs_per_step_per_gpu'], d['gpu_launches'], p and ('epe mean %.3g max %.3g' % (p['epe_gpu_vs_cpu_mean_px'], p['epe_gpu_vs_cpu_max_px'])), d['roofline']['stage_ms_per_step'])"
}
: > gpurun_out/extra_r01s.jsonl
echo "== C1 red-black"; run --sor redblack --steps 10 --cpu-seconds 4
echo "== C1 batch 1"; run --batch 1 --steps 200 --no-cpu-baseline
echo "== C1 batch 8"; run --batch 8 --steps 100 --no-cpu-baseline
echo "== C1 batch 148"; run --batch 148 --steps 50 --no-cpu-baseline
echo "== C4 batch 296"; run --workload C4 --batch 296 --steps 2 --no-cpu-baseline
