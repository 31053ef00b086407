#!/bin/bash
# round 2, call v (1 GPU): final validation with the team SOR and the new automatic fused/split rule: whole GPU suite, smoke, default bench, reference arm
O=gpurun_out
echo "== no-alloc alone"; timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q 2>&1 | tail -4
echo "== full suite"; timeout 2400 python -m pytest tests -m gpu -x -q 2>&1 | tail -6
echo "== smoke"; timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
echo "== bench default"; timeout 900 python bench.py --steps 20 --warmup 3 > $O/r02v_bench_C1.json 2> $O/r02v_bench_C1.err; tail -2 $O/r02v_bench_C1.err
python - <<'PY'
import json
d=json.loads([l for l in open("gpurun_out/r02v_bench_C1.json") if l.startswith("{")][0])
print(d["value"], d["ms_per_step"], d["e2e"]["value"], d["e2e_ofclass_cut"]["value"], d["gpu_launches"], d["cpu_baseline"]["value"], d["parity"]["bit_identical"], d["clocks"])
print({k: (v["frac"], v["issue_active_frac"], v["traffic"]) for k, v in d["roofline"]["per_stage"].items()})
PY
echo "== reference arm"; timeout 900 python bench.py --impl reference --steps 20 --warmup 3 > $O/r02v_ref_C1.json 2>&1; tail -c 600 $O/r02v_ref_C1.json
