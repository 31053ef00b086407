#!/bin/bash
# round 2, call aa (1 GPU): staged taps (with the unchanged-window skip) for every grid step: op1 staged against direct loads; then
# the final validation: whole GPU suite, smoke, default bench, reference arm
O=gpurun_out
run() {
  timeout 600 python bench.py "$@" --warmup 3 --steps 10 --no-e2e --no-cpu-baseline 2>/dev/null | grep '^{' | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l); p=d.get('parity') or {}; print(round(d['value'],1), 'pairs/s', round(d['ms_per_step'],3), 'ms', d['config']['pairs_per_step_per_gpu'], 'search ms', d['roofline']['stage_ms_per_step']['inverse_search'], 'bit-identical', p.get('bit_identical'))"
}
: > $O/r02aa_op1.txt
for pm in 0 1 0 1; do
  echo -n "[patch_map=$pm] op1: " | tee -a $O/r02aa_op1.txt
  run --workload op1 --batch 1776 --patch-map $pm | tee -a $O/r02aa_op1.txt
done
echo "== full suite"; timeout 2400 python -m pytest tests -m gpu -x -q 2>&1 | tail -6
echo "== smoke"; timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
echo "== bench default"; timeout 900 python bench.py --steps 20 --warmup 3 > $O/r02aa_bench_C1.json 2> $O/r02aa_bench_C1.err; tail -2 $O/r02aa_bench_C1.err
python - <<'PY'
import json
d=json.loads([l for l in open("gpurun_out/r02aa_bench_C1.json") if l.startswith("{")][0])
print(d["value"], d["ms_per_step"], d["e2e"]["value"], d["e2e_ofclass_cut"]["value"], d["gpu_launches"], d["cpu_baseline"]["value"], d["parity"]["bit_identical"], d["clocks"])
print(d["roofline"]["stage_ms_per_step"])
PY
echo "== reference arm"; timeout 900 python bench.py --impl reference --steps 20 --warmup 3 > $O/r02aa_ref_C1.json 2>&1; tail -c 400 $O/r02aa_ref_C1.json
