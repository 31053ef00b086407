#!/bin/bash
# round 2, call x (1 GPU): every bench workload again with the final build (staged search by grid step, reciprocal / packed 12x12
# search, team SOR, new fused / split rule): GPU numbers + parity objects; the CPU arm of r02k stands (the reference did not change)
O=gpurun_out
: > $O/r02x_workloads.jsonl
run() {
  timeout 900 python bench.py "$@" --warmup 3 --no-e2e --no-cpu-baseline 2>/dev/null | grep '^{' | tee -a $O/r02x_workloads.jsonl | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l); p=d.get('parity')
    print(d['config']['workload'][:40], '|', round(d['value'],1), 'pairs/s', round(d['ms_per_step'],3), 'ms', d['config']['pairs_per_step_per_gpu'], '| parity', p and ('%.3g / %.3g %s' % (p['epe_gpu_vs_cpu_mean_px'], p['epe_gpu_vs_cpu_max_px'], p['bit_identical'])), d['roofline']['stage_ms_per_step'])"
}
run --workload op1 --steps 10
run --workload rgb2 --batch 1184 --steps 10
run --workload C3 --steps 10
run --workload C5 --steps 10
run --workload C5 --batch 592 --steps 10 --no-parity
run --workload op3 --batch 592 --steps 3
run --workload C4 --batch 148 --steps 2
run --workload op4 --batch 148 --steps 2
run --workload C2 --batch 64 --steps 2
for b in 1 8 148; do run --batch $b --streams 1 --steps 20 --no-parity; done
run --workload C4 --batch 1 --steps 3 --no-parity
run --workload C4h --batch 1 --steps 3 --no-parity
