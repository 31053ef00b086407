#!/bin/bash
# round 2, call k (1 GPU): every bench workload with the round-2 CPU arm (reference core + cv2 pre/post) and its parity object
O=gpurun_out
: > $O/r02k_workloads.jsonl
run() {
  timeout 900 python bench.py "$@" --warmup 3 --no-e2e 2>/dev/null | grep '^{' | tee -a $O/r02k_workloads.jsonl | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l); p=d.get('parity'); c=d.get('cpu_baseline')
    print(d['config']['workload'][:40], '|', round(d['value'],1), 'pairs/s', d['config']['pairs_per_step_per_gpu'], '| cpu', c and (round(c['value'],2), round(c['core_pairs_per_s'],2), round(c['core_ms_1t'],2), round(c['prepost_ms_1t'],2)), '| parity', p and ('%.3g / %.3g %s' % (p['epe_gpu_vs_cpu_mean_px'], p['epe_gpu_vs_cpu_max_px'], p['bit_identical'])))"
}
run --workload op1 --steps 10 --cpu-seconds 20
run --workload rgb2 --batch 1184 --steps 10 --cpu-seconds 20
run --workload C3 --steps 10 --cpu-seconds 20
run --workload C5 --steps 10 --cpu-seconds 30
run --workload C5 --batch 592 --steps 10 --no-cpu-baseline --no-parity
run --workload op3 --batch 592 --steps 3 --cpu-seconds 60
run --workload C4 --batch 148 --steps 2 --cpu-seconds 60
run --workload op4 --batch 148 --steps 2 --cpu-seconds 150
run --workload C2 --batch 64 --steps 2 --cpu-seconds 300
