#!/usr/bin/env python
"""Raw pinned-memory PCIe bandwidth of this box with 1..N GPUs copying at the same time (one process per GPU).

  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29511 tools/pcie_probe.py

Plain cudaMemcpyAsync (torch .copy_ on a side stream) of 64 MB blocks between pinned host memory and the device, one copy
per block, for D2H alone, H2D alone and both directions together; all ranks start together (barrier) and run for the same
number of blocks; the aggregate GB/s is what bounds bench.py's `e2e` figure on this box (3.57 MB of f32 flow per C1 pair
must cross to the host).  Rank 0 prints one JSON line.
"""
import json
import os
import sys
import time

import torch
import torch.distributed as dist


def main():
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    affinity = "spread" if "--affinity-spread" in sys.argv else "none"
    if affinity == "spread":       # each rank on its own slice of the host cores BEFORE it allocates (first-touch placement)
        ncpu = os.cpu_count() or 1
        per = max(1, ncpu // world)
        try:
            os.sched_setaffinity(0, set(range(rank * per, min(ncpu, (rank + 1) * per))))
        except OSError:
            affinity = "spread (sched_setaffinity refused)"
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    blk = 64 << 20
    nblk = 48
    host = [torch.empty(blk, dtype=torch.uint8).pin_memory() for _ in range(4)]
    dev = [torch.empty(blk, dtype=torch.uint8, device="cuda") for _ in range(4)]
    s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
    out = {}

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for name in ("d2h", "h2d", "both"):
        for rep in range(2):           # the first repetition warms up
            barrier()
            t = time.perf_counter()
            for i in range(nblk):
                if name in ("d2h", "both"):
                    with torch.cuda.stream(s1):
                        host[i % 2].copy_(dev[i % 2], non_blocking=True)
                if name in ("h2d", "both"):
                    with torch.cuda.stream(s2):
                        dev[2 + i % 2].copy_(host[2 + i % 2], non_blocking=True)
            torch.cuda.synchronize()
            dt = time.perf_counter() - t
        gbs = blk * nblk * (2 if name == "both" else 1) / dt / 1e9
        v = torch.tensor([gbs, dt], device="cuda", dtype=torch.float64)
        if world > 1:
            lst = [torch.zeros_like(v) for _ in range(world)]
            dist.all_gather(lst, v)
        else:
            lst = [v]
        per = [float(x[0]) for x in lst]
        slowest = max(float(x[1]) for x in lst)
        out[name] = {"per_gpu_gbs": [round(x, 2) for x in per], "aggregate_gbs": round(blk * nblk * (2 if name == "both" else 1) * world / slowest / 1e9, 2)}
    if rank == 0:
        print(json.dumps({"probe": "pinned cudaMemcpyAsync, 64 MB blocks, %d blocks per direction" % nblk, "n_gpus": world, "cpu_count": os.cpu_count(), "affinity": affinity, **out}))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
