#!/usr/bin/env python
"""Latency of ONE 3840x2160 frame pair tiled spatially over 1..N GPUs of this box.

  gpurun --gpus N -- python tools/tiled_latency.py [--out profiles/r02_tiled_latency_N.json]

For both C4 parameter lines of SURVEY 8(d) (light: 8x8 patches down to level 2; heavy: 12x12 patches at overlap .75 down
to level 1), forward-backward consistency + L1 cost: device time per pair (disb200_tiled_time: mean of `reps` runs between
two all-device barriers, frames resident), end-to-end time through disb200_tiled_run_host (frames from host memory, result
to host memory), the level plan, and whether the result equals the one-device handle bit for bit.
"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import of_dis_b200 as dis  # noqa: E402
from of_dis_b200.synth import make_pair  # noqa: E402

LINES = {"light": dict(sc_f=7, sc_l=2, max_iter=16, min_iter=16, patchsz=8, patove=0.4, usefbcon=1, costfct=1),
         "heavy": dict(sc_f=7, sc_l=1, max_iter=16, min_iter=16, patchsz=12, patove=0.75, usefbcon=1, costfct=1)}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default=None)
    ap.add_argument("--reps", type=int, default=5)
    ap.add_argument("--virtual", type=int, default=0, help="also time K strips on ONE device (device 0 named K times)")
    args = ap.parse_args()
    ngpu = dis.device_count()
    w0, h0 = 3840, 2160
    A, B, _ = make_pair(h0, w0, seed=0)
    res = {"frame": [w0, h0, 1], "gpus_on_box": ngpu, "reps": args.reps, "lines": {}}
    for name, line in LINES.items():
        p = dis.operating_point(2, w0).copy(verbosity=0, **line)
        with dis.Handle(p, w0, h0, max_batch=1) as hd:
            want = hd.run_host(A[None], B[None])[0]
        rows = []
        sets = [list(range(n)) for n in (1, 2, 4, 8) if n <= ngpu]
        if args.virtual > 1:
            sets.append([0] * args.virtual)
        for devs in sets:
            with dis.Tiled(p, w0, h0, devs) as td:
                got = td.run(A, B)
                t = time.perf_counter()
                td.run(A, B)
                e2e_ms = (time.perf_counter() - t) * 1e3
                dev_ms = td.time_ms(args.reps)
                plan = {l: td.plan(l) for l in range(p.sc_l, p.sc_f + 1)}
                rows.append({"devices": devs, "device_ms_per_pair": dev_ms, "host_to_host_ms_per_pair": e2e_ms, "launches": td.launch_count(),
                             "bit_identical_to_one_device": bool(np.array_equal(got, want)),
                             "tiled_levels": [l for l, pl in plan.items() if pl[0]],
                             "strip_rows_finest": plan[p.sc_l][1]})
            print(name, rows[-1], flush=True)
        res["lines"][name] = {"params": " ".join("%g" % v for v in p.cli_args()), "runs": rows}
    if args.out:
        with open(args.out, "w") as f:
            json.dump(res, f, indent=1)
    print(json.dumps(res))


if __name__ == "__main__":
    main()
