#!/usr/bin/env python
"""bench.py - DIS optical flow hot path, frame pairs/s on B200 (contract: see DESIGN.md "Measurement").

  python bench.py [--gpus N] [--steps K] [--warmup W] [--batch B] [--workload C1|C2|C3|op1..op4]
  python -m torch.distributed.run --nproc-per-node N ... bench.py --gpus N ...
  python bench.py --impl reference ...      # the reference's CPU code on the host cores

A step is one pass of the whole per-pair path (pyramid -> inverse search -> densification ->
variational refinement -> upsample) over one batch of B synthetic frame pairs per GPU.
`value` is device-resident throughput (frames already in HBM, result left in HBM), `e2e` is the
same metric through disb200_run_batch_host with pinned HOST buffers (H2D + D2H inside the timed
region).  Prints ONE JSON line on rank 0.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WORKLOADS = {
    # name: (mode, channels, op, w0, h0, description)
    "C1": (1, 1, 2, 1024, 436, "run_OF_INT op2 1024x436 grey"),
    "C2": (1, 3, 4, 1024, 436, "run_OF_RGB op4 1024x436 BGR"),
    "C3": (2, 1, 2, 1280, 720, "run_DE_INT op2 1280x720 grey stereo"),
    "rgb2": (1, 3, 2, 1024, 436, "run_OF_RGB op2 1024x436 BGR"),
    "op1": (1, 1, 1, 1024, 436, "run_OF_INT op1 1024x436 grey"),
    "op3": (1, 1, 3, 1024, 436, "run_OF_INT op3 1024x436 grey"),
    "op4": (1, 1, 4, 1024, 436, "run_OF_INT op4 1024x436 grey"),
    "C5": (1, 1, 2, 1920, 1080, "run_OF_INT op2 1920x1080 grey video pairs"),
    # SURVEY.md 8(d) C4, lighter explicit line: fwd/bwd consistency, L1 cost, finest level 960x544
    "C4": (1, 1, 2, 3840, 2160, "run_OF_INT explicit params 3840x2160 grey, fb consistency + L1"),
    "C4h": (1, 1, 2, 3840, 2160, "run_OF_INT explicit params 3840x2160 grey, fb consistency + L1, 12x12 patches to level 1"),
}
# BASELINE.json does not pin C4's 20 parameters.  "C4" is the survey's lighter line (87 k patches, finest
# level 960x544), "C4h" its proposed heavy one (12x12 patches at overlap .75 down to level 1: 620 k patches fwd + bwd).
C4_LINES = {"C4": dict(sc_f=7, sc_l=2, max_iter=16, min_iter=16, patchsz=8, patove=0.4, usefbcon=1, costfct=1),
            "C4h": dict(sc_f=7, sc_l=1, max_iter=16, min_iter=16, patchsz=12, patove=0.75, usefbcon=1, costfct=1)}
N_DISTINCT = 4  # distinct synthetic pairs, tiled over the batch
DEFAULT_BATCH = 1776  # 2 sub-batch streams x 6 resident one-warp SORs x 148 SMs


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


def make_inputs(mode, ch, w0, h0, n_distinct=N_DISTINCT):
    from of_dis_b200.synth import make_pair
    As, Bs, Gs = [], [], []
    for s in range(n_distinct):
        A, B, gt = make_pair(h0, w0, channels=ch, mode="OF" if mode == 1 else "DE", seed=s)
        As.append(A)
        Bs.append(B)
        Gs.append(gt)
    return np.stack(As), np.stack(Bs), np.stack(Gs)


class ClockSampler(threading.Thread):
    """Samples SM clock and throttle reasons through NVML while the timed region runs."""

    def __init__(self, device):
        super().__init__(daemon=True)
        self.device, self.samples, self.reasons, self.stop_flag, self.max_mhz = device, [], set(), False, None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(device)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def run(self):
        if self.nv is None:
            return
        nv = self.nv
        names = {
            nv.nvmlClocksEventReasonHwSlowdown if hasattr(nv, "nvmlClocksEventReasonHwSlowdown") else 0x8: "hw_slowdown",
            getattr(nv, "nvmlClocksEventReasonHwThermalSlowdown", 0x40): "hw_thermal_slowdown",
            getattr(nv, "nvmlClocksEventReasonSwThermalSlowdown", 0x20): "sw_thermal_slowdown",
            getattr(nv, "nvmlClocksEventReasonSwPowerCap", 0x4): "sw_power_cap",
            getattr(nv, "nvmlClocksEventReasonHwPowerBrakeSlowdown", 0x80): "hw_power_brake",
        }
        while not self.stop_flag:
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                try:
                    r = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                except Exception:
                    r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for bit, name in names.items():
                    if r & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            time.sleep(0.05)

    def result(self):
        self.stop_flag = True
        self.join(timeout=1.0)
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": ["nvml unavailable"]}
        return {"sm_mhz": statistics.median(self.samples), "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons)}


def to_orc(p):
    from oracle.bindings import OrcParams
    q = OrcParams()
    for f, _ in OrcParams._fields_:
        setattr(q, f, getattr(p, {"psz": "patchsz", "noc": "channels"}.get(f, f)))
    return q


def orc_workload_params(workload):
    """The workload's parameters as the oracle's struct, built WITHOUT touching libdisb200.so (the reference arm
    must not load the product library)."""
    from oracle import bindings as ob
    mode, ch, op, w0, h0, _ = WORKLOADS[workload]
    q = ob.Port().operating_point(op, w0, mode, ch).copy(verbosity=0)
    if workload in C4_LINES:
        q = q.copy(**{{"patchsz": "psz"}.get(k, k): v for k, v in C4_LINES[workload].items()})
    return q


def repo_libs_loaded():
    """Shared objects under the repository mapped into this process (the reference arm must list oracle/ only)."""
    libs = set()
    try:
        with open("/proc/self/maps") as f:
            for ln in f:
                path = ln.split()[-1]
                if path.startswith(ROOT) and ".so" in path:
                    libs.add(os.path.relpath(path, ROOT))
    except OSError:
        pass
    return sorted(libs)


def orc_cli_args(q):
    return [q.sc_f, q.sc_l, q.max_iter, q.min_iter, q.dp_thresh, q.dr_thresh, q.res_thresh, q.psz, q.patove, q.usefbcon,
            q.patnorm, q.costfct, q.usetvref, q.tv_alpha, q.tv_gamma, q.tv_delta, q.tv_innerit, q.tv_solverit, q.tv_sor, q.verbosity]


class CpuArm:
    """The reference's CPU implementation of the path, split the way SURVEY 8(d) asks for it:

      core     = the OFC::OFClass constructor (oflow.cpp:31-338, the reference's own "O.Flow Run-Time"), from
                 oracle/_ref (the reference's sources compiled unmodified) when that library travelled with the
                 snapshot (kind "reference"), else the restatement (kind "port")
      pre/post = dimension padding + ConstructImgPyramide for both frames + resize/crop (run_dense.cpp:131-179,
                 299-312,407-415): python-cv2 (the same OpenCV the reference links, SIMD build; kind_prepost "cv2"),
                 else the scalar restatement ("port")
    """

    def __init__(self, q, A, B, force_port_prepost=False):
        from oracle import bindings as ob
        from oracle import cv2_prepost as cp
        self.q, self.A, self.B = q, A, B
        self.port = ob.Port()
        self.kind, self.ref = "port", None
        if ob.have_ref(q.mode, q.noc, bool(q.gradmag)):
            try:
                self.ref = ob.Ref(q.mode, q.noc, bool(q.gradmag))
                self.kind = "reference"
            except Exception:
                self.ref = None
        self.cp = cp if (cp.HAVE_CV2 and not force_port_prepost) else None
        self.kind_prepost = "cv2" if self.cp else "port"
        self.h0, self.w0 = A.shape[1:3]
        self.W, self.H, self.padl, self.padt = self.port.pad_dims(self.w0, self.h0, q.sc_f)

    def pre(self, i):
        a, b = self.A[i % len(self.A)], self.B[i % len(self.B)]
        q = self.q
        if self.cp:
            return (self.cp.pyramid(self.cp.pad_frame(a, q.sc_f)[0], q.sc_f, q.psz, bool(q.gradmag)),
                    self.cp.pyramid(self.cp.pad_frame(b, q.sc_f)[0], q.sc_f, q.psz, bool(q.gradmag)))
        return (self.port.pyramid(a, q.sc_f, q.sc_l, q.psz, bool(q.gradmag)), self.port.pyramid(b, q.sc_f, q.sc_l, q.psz, bool(q.gradmag)))

    def core(self, PA, PB):
        return (self.ref or self.port).ofclass(self.q, PA, PB, self.W, self.H)

    def post(self, low):
        if self.cp:
            return self.cp.upsample_crop(low, self.q.sc_l, self.padl, self.padt, self.w0, self.h0)
        return self.port.upsample_crop(low, self.q.sc_l, self.padl, self.padt, self.w0, self.h0)

    def one(self, i):
        return self.post(self.core(*self.pre(i)))

    def measure(self, seconds_budget, workers, windows=1, warm_windows=0):
        """-> dict.  Single-threaded worker PROCESSES ("N independent pairs on N host cores, one
        process-level worker per core"), all counting the pairs they finish inside common fixed windows: first ONE
        worker (the reference's default build has no OpenMP: per-pair ms of the core and of pre/post), then `workers`
        of them = the whole host.  `value` = core + pre/post on all cores, median over `windows` windows."""
        w1 = max(1.0, seconds_budget * 0.08)
        a1, c1, ra1, rc1 = run_cpu_workers(self, 1, w1, 1)
        core1, all1 = 1.0 / rc1, 1.0 / ra1
        nwin = windows + warm_windows
        wN = max(0.5, seconds_budget * 0.3 / nwin)
        aN, cN, raN, rcN = run_cpu_workers(self, workers, wN, nwin)
        if nwin == 1:      # one window: the workers' own rates (exact also when a pair takes longer than the window)
            v_all, v_core = raN, rcN
        else:
            v_all = statistics.median(aN[warm_windows:]) / wN
            v_core = statistics.median(cN[warm_windows:]) / wN
        return {"value": v_all, "unit": "pairs/s", "cores": workers, "kind": self.kind, "kind_prepost": self.kind_prepost,
                "core_ms_1t": core1 * 1e3, "prepost_ms_1t": (all1 - core1) * 1e3, "pairs_per_s_1t": 1.0 / all1, "core_pairs_per_s_1t": 1.0 / core1,
                "core_pairs_per_s": v_core, "core_plus_prepost_pairs_per_s": v_all,
                "per_window_pairs_per_s": [round(x / wN, 1) for x in aN],
                "sample": "%d single-threaded worker processes, %d windows of %.2f s each for (core + pre/post) and for (core only), median "
                          "of the last %d: %d and %d pairs per window; one worker alone: %d and %d pairs in %.1f s"
                          % (workers, nwin, wN, windows, statistics.median(aN[warm_windows:]), statistics.median(cN[warm_windows:]), a1[0], c1[0], w1)}


def run_cpu_workers(arm, workers, window, nwin):
    """Spawn `workers` fresh single-threaded processes of this script (--cpu-worker); all start at the same wall-clock
    instant and count the pairs they finish inside `nwin` consecutive `window`-second slots, first of the whole step
    (pyramids + OFClass + resize), then of OFClass alone.  -> (per-window totals whole step, per-window totals core)."""
    import subprocess
    import tempfile
    with tempfile.TemporaryDirectory() as td:
        npz = os.path.join(td, "in.npz")
        np.savez(npz, A=arm.A, B=arm.B, q=np.frombuffer(bytes(arm.q), np.uint8), port_prepost=np.array(arm.cp is None))
        start = time.time() + 4.0 + 0.05 * workers       # interpreter + library start-up of the workers
        env = dict(os.environ, OMP_NUM_THREADS="1", OPENBLAS_NUM_THREADS="1",
                   # steady-state allocator: keep the multi-MB image buffers on the heap instead of mmap/munmap + page
                   # faults per call - what a long-running host process sees (measured: 326 -> 635 pairs/s on 8 cores)
                   MALLOC_MMAP_THRESHOLD_="33554432", MALLOC_TRIM_THRESHOLD_="1073741824", MALLOC_TOP_PAD_="268435456")
        procs = [subprocess.Popen([sys.executable, os.path.abspath(__file__), "--cpu-worker", npz, repr(start), repr(window), str(nwin)],
                                  stdout=subprocess.PIPE, env=env) for _ in range(workers)]
        n_all, n_core = [0] * nwin, [0] * nwin
        rate_all = rate_core = 0.0
        for pr in procs:
            out = pr.communicate()[0].decode().strip().splitlines()
            if pr.returncode != 0 or not out:
                raise RuntimeError("CPU baseline worker failed")
            r = json.loads(out[-1])
            if r["late"]:
                raise RuntimeError("CPU baseline worker started after the common start time; raise the start-up margin")
            n_all = [x + y for x, y in zip(n_all, r["n_all"])]
            n_core = [x + y for x, y in zip(n_core, r["n_core"])]
            rate_all += r["n_all_rate"]
            rate_core += r["n_core_rate"]
    return n_all, n_core, rate_all, rate_core


def cpu_worker_main(npz, start, window, nwin):
    from oracle.bindings import OrcParams
    d = np.load(npz)
    q = OrcParams.from_buffer_copy(d["q"].tobytes())
    arm = CpuArm(q, d["A"], d["B"], force_port_prepost=bool(d["port_prepost"]))
    pyrs = [arm.pre(i) for i in range(len(arm.A))]
    arm.one(0)                                           # warm-up
    late = time.time() > start
    res = {}
    i = 0
    for name, fn in (("n_all", arm.one), ("n_core", lambda k: arm.core(*pyrs[k % len(pyrs)]))):
        while time.time() < start:
            pass
        cnt = [0] * nwin
        end = start + nwin * window
        done, t = 0, start
        while True:
            fn(i); i += 1
            t = time.time()
            done += 1
            if t < end:
                cnt[int((t - start) / window)] += 1      # a pair counts for the window it finishes in
            elif done >= 2:                              # slow workloads (seconds per pair): at least two pairs, whatever the window
                break
        res[name] = cnt
        res[name + "_rate"] = done / (t - start)         # pairs/s of this worker over its whole phase
        start = max(end, t) + 0.5
    res["late"] = late
    print(json.dumps(res))
    return 0


def cpu_runner(p, A, B):
    """(one, kind): one(i) -> full-resolution flow of pair i % len(A), the PARITY checker: the reference's OFClass
    (oracle/_ref when present) between the restatement's pyramid and upsample (bit-identical to cv2's on u8 frames,
    tests/test_oracle_cpu.py; cv2's own resize differs from it by <= 1e-5 px in the last step)."""
    arm = CpuArm(to_orc(p), A, B, force_port_prepost=True)
    return arm.one, arm.kind


def epe_stats(a, b):
    d = np.sqrt(((np.asarray(a, np.float64) - np.asarray(b, np.float64)) ** 2).sum(-1))
    return float(d.mean()), float(d.max())


def parity_report(p, A, B, G, hd):
    """north_star: EPE(GPU, CPU reference) mean / max and EPE against the synthetic ground truth, on the
    distinct pairs of the workload (the CPU flows come from the same code the cpu_baseline leg times)."""
    one, kind = cpu_runner(p, A, B)
    k = len(A)
    gpu = hd.run_host(np.ascontiguousarray(A), np.ascontiguousarray(B))
    cpu = np.stack([one(i).reshape(gpu.shape[1:]) for i in range(k)])
    m, mx = epe_stats(gpu, cpu)
    rep = {"pairs": k, "against": kind, "epe_gpu_vs_cpu_mean_px": m, "epe_gpu_vs_cpu_max_px": mx,
           "bit_identical": bool(np.array_equal(gpu, cpu)), "tolerance_mean_px": 0.01}
    if G is not None:
        rep["epe_vs_ground_truth_gpu_px"] = epe_stats(gpu, G)[0]
        rep["epe_vs_ground_truth_cpu_px"] = epe_stats(cpu, G)[0]
    return rep


def main():
    if len(sys.argv) == 6 and sys.argv[1] == "--cpu-worker":
        return cpu_worker_main(sys.argv[2], float(sys.argv[3]), float(sys.argv[4]), int(sys.argv[5]))
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--batch", type=int, default=None, help="frame pairs per step per GPU (default 1776 = two sub-batches of 6 resident SOR warps on each of the 148 SMs; C4: 16, C5: 256)")
    ap.add_argument("--workload", default="C1", choices=sorted(WORKLOADS))
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--cpu-seconds", type=float, default=20.0, help="CPU baseline sample budget (seconds of measuring windows)")
    ap.add_argument("--ref-seconds", type=float, default=150.0, help="--impl reference: total budget of the measuring windows")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--streams", type=int, default=0, help="DisTuning.streams (0 = automatic); 1 for a serialised ncu launch list")
    ap.add_argument("--tv-schedule", type=int, default=0, help="DisTuning.tv_schedule (0 auto, 1 fused, 2 split, 3 general wavefront)")
    ap.add_argument("--antiphase", type=int, default=0, help="DisTuning.antiphase (two sub-batches, refinements in anti-phase)")
    ap.add_argument("--set", default="", help="kernel experiments only: DisParams overrides k=v[,k=v] for the GPU arm (needs --no-parity "
                    "--no-cpu-baseline --no-e2e; the line's config names them)")
    ap.add_argument("--patch-map", type=int, default=0, help="DisTuning.patch_map (A/B of the inverse-search lane mappings)")
    ap.add_argument("--no-parity", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--sor", default="lexicographic", choices=["lexicographic", "redblack"],
                    help="redblack: the throughput variant (a different iterate: ~0.04 px mean from the parity result, never the headline)")
    args = ap.parse_args()
    W_ = max(args.warmup, 3) if args.impl == "ours" else args.warmup

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    mode, ch, op, w0, h0, desc = WORKLOADS[args.workload]
    nop = 2 if mode == 1 else 1

    q_orc = orc_workload_params(args.workload)     # oracle-side struct: the reference arm never loads libdisb200.so
    if args.workload in C4_LINES and args.batch is None:
        args.batch = 148 if args.workload == "C4" else 8   # C4: one CTA per SM in the general wavefront SOR; 570 pairs/s at 16, 1305 at 96
    seq = args.workload == "C5"       # video: consecutive frames, pair i = (i, i+1), pyramids shared
    if seq:
        from of_dis_b200.synth import make_sequence
        fr = make_sequence(h0, w0, 5, seed=0)
        A, B, G = np.stack(fr[:-1]), np.stack(fr[1:]), None
        if args.batch is None:
            args.batch = 256          # BASELINE.json configs[4]: 256 consecutive pairs
    else:
        A, B, G = make_inputs(mode, ch, w0, h0, 1 if args.workload in C4_LINES else N_DISTINCT)
    if args.batch is None:
        args.batch = DEFAULT_BATCH
    bytes_per_pair = 2 * w0 * h0 * ch + 4 * w0 * h0 * nop     # compulsory traffic of the whole path
    config = {"workload": "%s: %s, params %s" % (args.workload, desc, " ".join("%g" % v for v in orc_cli_args(q_orc))),
              "pairs_per_step_per_gpu": args.batch, "frame": [w0, h0, ch], "l2_policy": "inputs+outputs per step (%.0f MB) exceed the 126 MB L2" % (args.batch * bytes_per_pair / 1e6),
              "sor": "lexicographic wavefront (parity mode, bit-exact with solver.c sor_coupled)" if args.sor == "lexicographic"
                     else "red-black (NOT the parity mode: a different iterate of the same system; `parity` below is against the lexicographic CPU reference)"}

    # ------------------------------------------------------------------ reference arm (CPU)
    if args.impl == "reference":
        if rank != 0:
            return 0
        threads = os.cpu_count() or 1
        arm = CpuArm(q_orc, A, B)
        # every "step" is one window of the worker processes' run: a bounded sample of the workload (whole run ~ 2 min)
        cpu = arm.measure(args.ref_seconds, threads, windows=max(args.steps, 1), warm_windows=max(args.warmup, 0))
        v = cpu["value"]
        config["pairs_per_step_per_gpu"] = None          # the CPU arm has no batch: a step is a bounded sample of pairs
        config["sample_per_step"] = cpu["sample"]
        line = {"impl": "reference", "metric": "frame_pairs_per_s", "value": v, "unit": "pairs/s", "mpix_per_s": v * w0 * h0 / 1e6,
                "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_pair_all_cores": 1e3 / v,
                "ms_per_step": 1e3 * args.batch / v, "ms_per_step_note": "time the host cores need for one %d-pair step of the GPU arm" % args.batch,
                "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                "config": config, "cpu_baseline": cpu, "native_so_loaded": repo_libs_loaded(),
                "e2e": {"value": v, "unit": "pairs/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
        print(json.dumps(line))
        return 0

    import of_dis_b200 as dis   # raises if libdisb200.so is missing: there is no fallback
    p = dis.operating_point(op, w0, mode, ch).copy(verbosity=0)
    if args.workload in C4_LINES:
        p = p.copy(**C4_LINES[args.workload])
    if args.sor == "redblack":
        p = p.copy(sor_mode=dis.SOR_REDBLACK)
    overrides = {}
    if args.set:
        if not (args.no_parity and args.no_cpu_baseline and args.no_e2e):
            ap.error("--set is for kernel experiments: pass --no-parity --no-cpu-baseline --no-e2e with it")
        for kv in args.set.split(","):
            k, v = kv.split("=")
            overrides[k] = float(v) if "." in v else int(v)
        p = p.copy(**overrides)
        config["param_overrides"] = overrides

    # ------------------------------------------------------------------ our arm (GPU)
    import torch
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the product has no CPU path")
    torch.cuda.set_device(local_rank)
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    nb = args.batch
    reps = -(-nb // len(A))
    hA = torch.from_numpy(np.ascontiguousarray(np.concatenate([A] * reps)[:nb])).pin_memory()
    hB = torch.from_numpy(np.ascontiguousarray(np.concatenate([B] * reps)[:nb])).pin_memory()
    hOut = torch.empty((nb, h0, w0, nop), dtype=torch.float32).pin_memory()
    dA, dB = hA.cuda(non_blocking=False), hB.cuda(non_blocking=False)
    dOut = torch.empty((nb, h0, w0, nop), dtype=torch.float32, device="cuda")
    torch.cuda.synchronize()
    if args.streams or args.tv_schedule or args.antiphase or args.patch_map:
        dis.set_tuning(streams=args.streams, tv_schedule=args.tv_schedule, antiphase=args.antiphase, patch_map=args.patch_map)
    hd = dis.Handle(p, w0, h0, max_batch=nb, device=local_rank)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    if seq:
        # ping-pong over the 5 synthetic frames so that every consecutive pair is a real, small motion
        order = [(i % 8) if (i % 8) < 5 else 8 - (i % 8) for i in range(nb + 1)]
        frames_np = np.stack([fr[k] for k in order])
        hF = torch.from_numpy(frames_np).pin_memory()
        dF = hF.cuda()

    def step():
        if seq:
            hd.run_sequence_device(dF.data_ptr(), nb + 1, dOut.data_ptr())
        else:
            hd.run_device(dA.data_ptr(), dB.data_ptr(), nb, dOut.data_ptr())

    for _ in range(W_):
        step()
    hd.sync()
    launches_per_step = hd.launch_count()
    sampler = ClockSampler(local_rank)
    sampler.start()
    barrier()
    hd.timer_begin()
    for _ in range(args.steps):
        step()
    ms = hd.timer_end()
    barrier()
    clocks = sampler.result()
    if world > 1:
        t = torch.tensor([ms], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    pairs = nb * args.steps * world
    value = pairs / (ms / 1e3)

    # per-stage device time (CUDA events on the launching stream) -> roofline of the dominant stage
    hd.set_stage_timing(True)
    samples = {}
    for _ in range(5):
        step()
        hd.sync()
        for k, v in hd.stage_ms().items():
            samples.setdefault(k, []).append(v)
    acc = {k: float(np.median(v)) for k, v in samples.items()}   # median of 5 passes: one host hiccup does not move a stage
    hd.set_stage_timing(False)
    fused_tv = acc.get("system", 0.0) == 0.0 and acc.get("sor", 0.0) > 0.0     # small batches: one kernel per level
    acc["varref_prep"] = acc.get("varref", 0.0) - acc.get("sor", 0.0) - acc.get("system", 0.0) - acc.get("smooth", 0.0)
    stage_share = {k: acc.get(k, 0.0) for k in ("pyramid", "inverse_search", "densify", "varref_prep", "smooth", "system", "sor", "upsample")}
    step_ms = sum(stage_share.values())
    peak, peak_src = peaks()
    # algorithmic bytes of each kernel group per pair
    W, H = hd.W, hd.H
    lv = [(W >> l, H >> l, l) for l in range(p.sc_l, p.sc_f + 1)]
    pad = p.patchsz
    C_ = ch
    steps_ = max(1, int(np.floor(np.float32(pad) * (np.float32(1) - np.float32(p.patove)))))
    npatch = [int(np.ceil(w / steps_)) * int(np.ceil(h / steps_)) for w, h, _ in lv]
    novals = C_ * pad * pad
    ndir = 2 if p.usefbcon else 1
    tv = 1 if p.usetvref else 0
    n_in = lambda l: p.tv_innerit * (l + 1)                       # fixed-point iterations at level l (refine_variational.cpp:36)
    alg = {
        # read both u8 frames, write image + dx + dy of every level for both frames
        "pyramid": 2 * w0 * h0 * ch + sum(2 * 3 * 4 * ch * (w + 2 * pad) * (h + 2 * pad) for w, h, _ in lv),
        # per level and direction: template image/dx/dy + target image once, per patch p_iter, pt_iter, weights out
        "inverse_search": ndir * sum(4 * 4 * ch * (w + 2 * pad) * (h + 2 * pad) + npatch[k] * 4 * (nop + 2 + novals)
                                     for k, (w, h, _) in enumerate(lv)),
        # patch results in, dense flow out
        "densify": ndir * sum(npatch[k] * 4 * (nop + novals) + 4 * nop * w * h for k, (w, h, _) in enumerate(lv)),
        # warp + derivatives once per level (2 images + flow in, 9C+1 planes out), finish: flow + increment in, flow out
        "varref_prep": tv * ndir * sum(4 * w * h * ((2 * ch + nop) + (9 * ch + 1) + 3 * nop) for w, h, l in lv),
        # per fixed-point iteration: flow + increment in, diffusivity {h, v} out
        "smooth": tv * ndir * sum(n_in(l) * 4 * w * h * (2 * nop + 2) for w, h, l in lv),
        # per fixed-point iteration: 8C derivative planes + mask + diffusivity + flow + increment in, two float4 packs out
        "system": tv * ndir * sum(n_in(l) * 4 * w * h * (8 * ch + 1 + 2 + 2 * nop + 8) for w, h, l in lv),
        "upsample": 4 * nop * lv[0][0] * lv[0][1] + 4 * nop * w0 * h0,
        # per fixed-point iteration the SOR makes ONE pass over memory whatever the number of sweeps (they are pipelined
        # inside the kernel): two float4 coefficient packs + the increment in, the increment out
        "sor": tv * ndir * sum(n_in(l) * 4 * w * h * (8 + 2 * nop) for w, h, l in lv),
    }
    nl = len(lv) * ndir
    n_it = ndir * sum(n_in(l) for _, _, l in lv)
    kernels = {"pyramid": "downscale_u8_kernel + halve_kernel + sobel_pad_kernel", "inverse_search": "patch_search_kernel (%d launches)" % nl,
               "densify": "densify_kernel (%d launches)" % nl, "varref_prep": "warp_kernel, deriv1_kernel, deriv2_kernel, finish_kernel (%d launches each)" % nl,
               "smooth": "smooth_kernel2 (%d launches)" % n_it, "system": "system_kernel (%d launches)" % n_it,
               "sor": ("tv_fused_kernel (%d launches: diffusivity + system + SOR of a level in one CTA per pair)" % nl) if fused_tv
                      else "sor_team_kernel / sor_warp_kernel / sor_band_kernel / sor_wavefront_kernel (%d launches)" % n_it,
               "upsample": "upsample_crop_kernel (1 launch)"}
    if fused_tv:
        alg["sor"] += alg["smooth"] + alg["system"]
    dom = max(stage_share, key=stage_share.get)
    dom_ms = stage_share[dom]
    dom_bytes = alg.get(dom)
    # DRAM traffic per pair and stage, and the issue-slot utilisation of each stage, from the committed ncu --set full
    # capture of the shipped kernels (profiles/r02_kernels.json, built by tools/ncu_summary.py; C1 only)
    traffic, issue = {}, {}
    tpath = os.path.join(ROOT, "profiles", "r02_kernels.json")
    if args.workload == "C1" and os.path.exists(tpath):
        with open(tpath) as f:
            ks = json.load(f)["per_stage"]
        traffic = {k: v["dram_read_bytes_per_pair"] + v["dram_write_bytes_per_pair"] for k, v in ks.items()}
        issue = {k: v.get("issue_active_pct") for k, v in ks.items()}
    roof = {"bound": "hbm", "kernel": kernels[dom], "stage": dom, "unit": "GB/s", "peak": peak, "peak_source": peak_src,
            "share_of_step": dom_ms / step_ms if step_ms else None,
            "traffic": traffic[dom] * nb if dom in traffic else None}
    if dom_bytes:
        roof["achieved"] = dom_bytes * nb / (dom_ms / 1e3) / 1e9
        roof["frac"] = roof["achieved"] / peak
        roof["algorithmic_bytes_per_launch_set"] = dom_bytes * nb
    else:
        roof["achieved"] = None
        roof["frac"] = None
    # every stage: algorithmic GB/s against the same peak (the launch set of the stage, per step)
    roof["per_stage"] = {k: {"ms": round(v, 4), "algorithmic_bytes": alg[k] * nb, "achieved": round(alg[k] * nb / (v / 1e3) / 1e9, 1) if v > 0 else None,
                             "frac": round(alg[k] * nb / (v / 1e3) / 1e9 / peak, 4) if v > 0 else None,
                             "traffic": traffic[k] * nb if k in traffic else None,
                             # issue-slot roofline (the one that bounds the latency / instruction-bound stages): share of the
                             # SM sub-partitions' issue cycles in which an instruction was issued, ncu smsp__issue_active
                             "issue_active_frac": round(issue[k] / 100.0, 4) if issue.get(k) is not None else None} for k, v in stage_share.items()}
    roof["issue_active_frac"] = roof["per_stage"][dom]["issue_active_frac"]
    path_gbs = value / world * bytes_per_pair / 1e9
    roof["whole_path"] = {"algorithmic_bytes_per_pair": bytes_per_pair, "achieved": path_gbs, "frac": path_gbs / peak}
    roof["stage_ms_per_step"] = {k: round(v, 4) for k, v in stage_share.items()}

    # ---- end to end through the host-pointer C-ABI entry (pinned host buffers) ----
    e2e = None

    def host_call():
        if seq:
            hd.run_sequence_host_ptr(hF.data_ptr(), nb + 1, hOut.data_ptr())
        else:
            hd.run_host_ptr(hA.data_ptr(), hB.data_ptr(), nb, hOut.data_ptr())

    if not args.no_e2e:
        for _ in range(2):
            host_call()
        barrier()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            host_call()
        barrier()
        dt = time.perf_counter() - t0
        if world > 1:
            t = torch.tensor([dt], device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            dt = float(t.item())
        e2e = {"value": pairs / dt, "unit": "pairs/s", "h2d_bytes_per_step": ((nb + 1) if seq else 2 * nb) * w0 * h0 * ch * world,
               "d2h_bytes_per_step": nb * w0 * h0 * nop * 4 * world, "checksum": float(hOut[0].double().abs().mean())}

    # The same call through the OFClass cut (u8 frames in, level-sc_l flow out: what OFC::OFClass writes to `outflow`,
    # oflow.h:105; the x 2^sc_l resize of run_dense.cpp:407-415 is left to the consumer).  Not the headline: it shows
    # what is left of the end-to-end cost once the full-resolution f32 flow (80 % of the compulsory bytes) does not
    # have to cross PCIe.
    e2e_low = None
    if not args.no_e2e:
        hl, wl, _ = hd.lowres_dims()
        hLow = torch.empty((nb, hl, wl, nop), dtype=torch.float32).pin_memory()

        def host_call_low():
            if seq:
                hd.run_sequence_host_lowres_ptr(hF.data_ptr(), nb + 1, hLow.data_ptr())
            else:
                hd.run_host_lowres_ptr(hA.data_ptr(), hB.data_ptr(), nb, hLow.data_ptr())

        for _ in range(2):
            host_call_low()
        barrier()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            host_call_low()
        barrier()
        dt = time.perf_counter() - t0
        if world > 1:
            t = torch.tensor([dt], device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            dt = float(t.item())
        e2e_low = {"value": pairs / dt, "unit": "pairs/s", "what": "u8 frames in, level-sc_l flow out (OFClass outflow, %dx%dx%d f32 per pair)" % (wl, hl, nop),
                   "h2d_bytes_per_step": e2e["h2d_bytes_per_step"], "d2h_bytes_per_step": nb * hl * wl * nop * 4 * world,
                   "checksum": float(hLow[0].double().abs().mean())}

    cpu = parity = None
    if rank == 0 and world == 1:
        if not args.no_cpu_baseline:
            cpu = CpuArm(q_orc, A, B).measure(args.cpu_seconds, os.cpu_count() or 1)
        if not args.no_parity:
            parity = parity_report(p, A, B, G, hd)

    if rank == 0:
        line = {"metric": "frame_pairs_per_s", "value": value, "unit": "pairs/s", "mpix_per_s": value * w0 * h0 / 1e6,
                "n_gpus": world, "steps": args.steps, "warmup": W_, "ms_per_step": ms / args.steps,
                "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
                "data": "synthetic warped-texture pairs (%d distinct seeds tiled over the batch)" % len(A),
                "config": config, "clocks": clocks, "e2e": e2e, "e2e_ofclass_cut": e2e_low, "gpu_launches": launches_per_step * args.steps,
                "roofline": roof, "cpu_baseline": cpu, "parity": parity}
        print(json.dumps(line))
    hd.close()
    if world > 1:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
