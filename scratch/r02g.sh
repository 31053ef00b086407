#!/bin/bash
# round 2, call g (8 GPUs): PCIe probe 1/2/4/8, bench at 8 GPUs (e2e), tiled over 8 GPUs
O=gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
nvidia-smi topo -m > $O/r02g_topo_8.txt 2>&1
lscpu | head -25 > $O/r02g_lscpu.txt 2>&1
: > $O/r02g_pcie_probe.jsonl
for n in 1 2 4 8; do
  $TR --nproc-per-node $n --master-port $((29600+n)) tools/pcie_probe.py 2>/dev/null | grep '^{' | tee -a $O/r02g_pcie_probe.jsonl
done
$TR --nproc-per-node 8 --master-port 29620 tools/pcie_probe.py --affinity-spread 2>/dev/null | grep '^{' | tee -a $O/r02g_pcie_probe.jsonl
echo "== bench 8 GPUs"
timeout 900 $TR --nproc-per-node 8 --master-port 29630 bench.py --gpus 8 --steps 10 --warmup 3 2>$O/r02g_bench8.err | grep '^{' > $O/r02g_bench_C1_8gpu.json; tail -2 $O/r02g_bench8.err
python - <<'PY'
import json
d=json.loads(open("gpurun_out/r02g_bench_C1_8gpu.json").read())
print(d["value"], d["ms_per_step"], d["e2e"], d["e2e_ofclass_cut"]["value"])
PY
echo "== tiled 8"; timeout 900 python -m pytest tests/test_tiled_multi_gpu.py -m gpu -q -k "physical" 2>&1 | tail -3
timeout 900 python tools/tiled_latency.py --out $O/r02g_tiled_latency_8.json --reps 5 2>&1 | grep -v "^{" | tail -10
