#!/bin/bash
# round 2, call b: full GPU suite on the widened parameter domain + new entry points; bench sanity; heavy C4 line split
mkdir -p gpurun_out
O=gpurun_out
echo "== full suite"; timeout 2400 python -m pytest tests -m gpu -x -q 2>&1 | tail -15
echo "== bench C1 (no e2e, no cpu)"; timeout 600 python bench.py --steps 10 --warmup 3 --no-e2e --no-cpu-baseline 2>$O/r02b_c1.err | grep '^{' > $O/r02b_bench_C1.json; tail -2 $O/r02b_c1.err
python - <<'PY'
import json
d=json.loads(open("gpurun_out/r02b_bench_C1.json").read())
print(d["value"], d["ms_per_step"], d["gpu_launches"], d["roofline"]["stage_ms_per_step"], d["parity"] and d["parity"]["bit_identical"])
PY
for wl in C4h C4; do
  for b in 1 4; do
    echo "== $wl batch $b"
    timeout 900 python bench.py --workload $wl --batch $b --steps 3 --warmup 3 --no-e2e --no-cpu-baseline $( [ $b != 1 ] && echo --no-parity ) 2>$O/r02b_$wl.err | grep '^{' > $O/r02b_bench_${wl}_b$b.json; tail -2 $O/r02b_$wl.err
    python - $wl $b <<'PY'
import json,sys
d=json.loads(open("gpurun_out/r02b_bench_%s_b%s.json"%(sys.argv[1],sys.argv[2])).read())
print(d["value"], "pairs/s", d["ms_per_step"], "ms/step", d["gpu_launches"], d["roofline"]["stage_ms_per_step"], d["parity"])
PY
  done
done
