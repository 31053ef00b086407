#!/bin/bash
# round 2, call a: sanity (suite + bench with the new CPU arm), staggered sub-batch sweep, tolerance build, sanitizer
mkdir -p gpurun_out
O=gpurun_out
echo "== full suite"; timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
echo "== bench default"; timeout 900 python bench.py --steps 20 --warmup 3 > $O/r02a_bench_C1.json 2> $O/r02a_bench_C1.err; tail -2 $O/r02a_bench_C1.err
python - <<'PY'
import json
d=json.loads([l for l in open("gpurun_out/r02a_bench_C1.json") if l.startswith("{")][0])
print(d["value"], d["ms_per_step"], d["e2e"]["value"], d["gpu_launches"], d["cpu_baseline"], d["parity"]["bit_identical"], d["clocks"])
PY
echo "== reference arm"; timeout 600 python bench.py --impl reference --steps 5 --warmup 1 --ref-seconds 60 > $O/r02a_ref_C1.json 2>&1; tail -c 1500 $O/r02a_ref_C1.json
run() {
  timeout 300 python bench.py "$@" --warmup 3 --no-e2e --no-cpu-baseline 2>/dev/null | grep '^{' | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l); print(round(d['value'],1), 'pairs/s', round(d['ms_per_step'],3), 'ms', d['config']['pairs_per_step_per_gpu'], d['gpu_launches'])"
}
: > $O/r02a_stagger.txt
for cfg in "2 0" "2 1" "2 9" "2 16" "2 17" "2 18" "2 19" "2 20" "2 21" "3 0" "3 17" "3 19" "3 20" "4 0" "4 17" "4 19" "1 0"; do
  set -- $cfg
  echo -n "streams $1 stagger $2: " | tee -a $O/r02a_stagger.txt
  DISB200_STREAMS=$1 DISB200_STAGGER=$2 run --steps 10 | tee -a $O/r02a_stagger.txt
done
echo "== batch 3552, streams 4 stagger 19"; DISB200_STREAMS=4 DISB200_STAGGER=19 run --steps 6 --batch 3552 | tee -a $O/r02a_stagger.txt
echo "== C5 256 streams 2 stagger 19"; DISB200_STREAMS=2 DISB200_STAGGER=19 run --steps 10 --workload C5 | tee -a $O/r02a_stagger.txt
echo "== C5 256 default"; run --steps 10 --workload C5 | tee -a $O/r02a_stagger.txt
echo "== tolerance build (FMA + approximate div/sqrt)"
DISB200_LIB=$PWD/of_dis_b200/libdisb200_tol.so timeout 600 python bench.py --steps 10 --warmup 3 --no-e2e --cpu-seconds 4 2>/dev/null | grep '^{' > $O/r02a_bench_C1_tol.json
python - <<'PY'
import json
d=json.loads(open("gpurun_out/r02a_bench_C1_tol.json").read())
print("tol", d["value"], d["roofline"]["stage_ms_per_step"], d["parity"])
d=json.loads([l for l in open("gpurun_out/r02a_bench_C1.json") if l.startswith("{")][0])
print("par", d["value"], d["roofline"]["stage_ms_per_step"])
PY
echo "== sanitizer memcheck"
timeout 900 compute-sanitizer --tool memcheck --log-file $O/r02a_memcheck.log python -m pytest tests/test_parity_gpu.py -x -q -k "tv_paths or band_sor or C1_of_int_op2 or varref_sweep or of_int_fb" 2>&1 | tail -2
tail -3 $O/r02a_memcheck.log
echo "== sanitizer initcheck"
timeout 600 compute-sanitizer --tool initcheck --log-file $O/r02a_initcheck.log python -m pytest tests/test_parity_gpu.py -x -q -k "varref_sweep or band_sor" 2>&1 | tail -2
tail -3 $O/r02a_initcheck.log
