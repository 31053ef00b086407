#!/bin/bash
# One GPU round: tests, bench, ncu launch list.  Usage: gpu_round.sh TAG [skip_tests]
TAG=${1:-r01g}
OUT=gpurun_out/$TAG
mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.max.sm,clocks.sm,power.limit --format=csv > $OUT/smi.txt 2>&1
if [ -z "$2" ]; then
  timeout 900 python -m pytest tests -m gpu -x -q > $OUT/pytest.log 2>&1; echo "pytest rc=$?" >> $OUT/pytest.log
  tail -3 $OUT/pytest.log
fi
timeout 600 python bench.py --steps 20 --warmup 3 > $OUT/bench.json 2> $OUT/bench.err; echo "bench rc=$?"
cat $OUT/bench.json | cut -c1-300
L=$(python -c "import json;d=json.load(open('$OUT/bench.json'));print(d['gpu_launches']//d['steps'])")
echo "launches per step: $L"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -s $((3*L)) -c $((2*L)) --csv --log-file $OUT/launches.csv \
  python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu-baseline > $OUT/ncu_launch.log 2>&1; echo "ncu list rc=$?"
ls -la $OUT
