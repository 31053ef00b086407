"""of_dis_b200 - Python host bindings of libdisb200.so (include/disb200.h).

The product is the CUDA library; this module is the thin ctypes layer the tests and bench.py
call it through.  It never computes anything itself and has no CPU path: if the shared
library is missing, or there is no CUDA device, the calls raise.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
# DISB200_LIB selects another build of the same C-ABI (the tolerance build libdisb200_tol.so, see build.py)
LIB_PATH = os.environ.get("DISB200_LIB") or os.path.join(HERE, "libdisb200.so")

MODE_OF, MODE_DE = 1, 2
SOR_LEXICOGRAPHIC, SOR_REDBLACK = 0, 1
STAGES = ("pyramid", "inverse_search", "densify", "varref", "upsample", "sor", "system", "smooth")


class DisError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("disb200 error %d: %s" % (code, msg))
        self.code = code


class DisParams(C.Structure):
    """Mirror of struct DisParams (run_dense.cpp:272-292 order + mode/channels/sor_mode)."""
    _fields_ = [
        ("sc_f", C.c_int), ("sc_l", C.c_int), ("max_iter", C.c_int), ("min_iter", C.c_int),
        ("dp_thresh", C.c_float), ("dr_thresh", C.c_float), ("res_thresh", C.c_float),
        ("patchsz", C.c_int), ("patove", C.c_float),
        ("usefbcon", C.c_int), ("patnorm", C.c_int), ("costfct", C.c_int), ("usetvref", C.c_int),
        ("tv_alpha", C.c_float), ("tv_gamma", C.c_float), ("tv_delta", C.c_float),
        ("tv_innerit", C.c_int), ("tv_solverit", C.c_int), ("tv_sor", C.c_float),
        ("verbosity", C.c_int), ("mode", C.c_int), ("channels", C.c_int), ("sor_mode", C.c_int), ("gradmag", C.c_int),
    ]

    def copy(self, **kw):
        q = DisParams()
        C.memmove(C.byref(q), C.byref(self), C.sizeof(DisParams))
        for k, v in kw.items():
            setattr(q, k, v)
        return q

    @property
    def nop(self):
        return 2 if self.mode == MODE_OF else 1

    def cli_args(self):
        """The 20 positional parameters in the reference's CLI order."""
        return [self.sc_f, self.sc_l, self.max_iter, self.min_iter, self.dp_thresh, self.dr_thresh, self.res_thresh,
                self.patchsz, self.patove, self.usefbcon, self.patnorm, self.costfct, self.usetvref, self.tv_alpha,
                self.tv_gamma, self.tv_delta, self.tv_innerit, self.tv_solverit, self.tv_sor, self.verbosity]


_lib = None


def lib():
    """Load libdisb200.so; raises if it has not been built (no fallback)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError("%s is missing: run `python -m of_dis_b200.build` (needs nvcc). "
                              "There is no CPU implementation to fall back to." % LIB_PATH)
        L = C.CDLL(LIB_PATH)
        L.disb200_last_error.restype = C.c_char_p
        L.disb200_version.restype = C.c_char_p
        L.disb200_free.restype = None
        L.disb200_writer_destroy.restype = None
        _lib = L
    return _lib


def _check(rc):
    if rc != 0:
        raise DisError(rc, lib().disb200_last_error().decode())


def _ptr(a, t=C.c_float):
    return None if a is None else a.ctypes.data_as(C.POINTER(t))


def operating_point(op: int, width: int, mode: int = MODE_OF, channels: int = 1) -> DisParams:
    p = DisParams()
    _check(lib().disb200_operating_point(op, width, mode, channels, C.byref(p)))
    return p


def padded_dims(w0: int, h0: int, sc_f: int):
    W, H, pl, pt = C.c_int(), C.c_int(), C.c_int(), C.c_int()
    _check(lib().disb200_padded_dims(w0, h0, sc_f, C.byref(W), C.byref(H), C.byref(pl), C.byref(pt)))
    return W.value, H.value, pl.value, pt.value


def device_count() -> int:
    return lib().disb200_device_count()


class Handle:
    """disb200_create / run_batch* / destroy."""

    def __init__(self, params: DisParams, w0: int, h0: int, max_batch: int = 1, device: int = 0):
        self.p = params.copy()
        self.w0, self.h0, self.max_batch, self.device = w0, h0, max_batch, device
        self.h = C.c_void_p()
        _check(lib().disb200_create(C.byref(self.p), w0, h0, max_batch, device, C.byref(self.h)))
        self.W, self.H, self.padl, self.padt = padded_dims(w0, h0, params.sc_f)

    def close(self):
        if self.h:
            lib().disb200_destroy(self.h)
            self.h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    # -- host buffers (numpy) ---------------------------------------------------------------
    def run_host(self, A: np.ndarray, B: np.ndarray, out: np.ndarray | None = None) -> np.ndarray:
        """A, B: uint8 (n, h0, w0[, C]) contiguous; returns float32 (n, h0, w0, nop)."""
        n = A.shape[0]
        assert A.dtype == np.uint8 and B.dtype == np.uint8 and A.flags.c_contiguous and B.flags.c_contiguous
        assert A.size == n * self.h0 * self.w0 * self.p.channels and B.shape == A.shape
        if out is None:
            out = np.empty((n, self.h0, self.w0, self.p.nop), np.float32)
        _check(lib().disb200_run_batch_host(self.h, _ptr(A, C.c_uint8), _ptr(B, C.c_uint8), n, _ptr(out)))
        return out

    def run_sequence(self, frames: np.ndarray, out: np.ndarray | None = None) -> np.ndarray:
        """frames: uint8 (nframes, h0, w0[, C]) contiguous -> float32 (nframes - 1, h0, w0, nop); pair i = (i, i+1)."""
        nf = frames.shape[0]
        assert frames.dtype == np.uint8 and frames.flags.c_contiguous and frames.size == nf * self.h0 * self.w0 * self.p.channels
        if out is None:
            out = np.empty((nf - 1, self.h0, self.w0, self.p.nop), np.float32)
        _check(lib().disb200_run_sequence_host(self.h, _ptr(frames, C.c_uint8), nf, _ptr(out)))
        return out

    def run_sequence_warm(self, frames: np.ndarray, out: np.ndarray | None = None) -> np.ndarray:
        """frames (n+1, h0, w0[, C]) u8 -> (n, h0, w0, nop) f32; pair i warm-started from pair i-1 (oflow.cpp:213-217)."""
        frames = np.ascontiguousarray(frames, np.uint8)
        n = frames.shape[0] - 1
        if out is None:
            out = np.empty((n, self.h0, self.w0, self.p.nop), np.float32)
        _check(lib().disb200_run_sequence_warm_host(self.h, _ptr(frames, C.c_uint8), n + 1, _ptr(out)))
        return out

    def ofclass(self, A, B, initflow=None) -> np.ndarray:
        """disb200_ofclass_handle: the OFClass cut on this handle's workspace (A, B as for ofclass())."""
        out = np.zeros((self.H >> self.p.sc_l, self.W >> self.p.sc_l, self.p.nop), np.float32)
        _check(lib().disb200_ofclass_handle(self.h, *_pyr_ptrs(self.p, A, B), _ptr(out), _ptr(initflow)))
        return out

    def lowres_dims(self):
        """(h_l, w_l, nop) of the level-sc_l flow that the *_lowres entry points return (OFClass's outflow, oflow.h:105)."""
        w, h, nop = C.c_int(), C.c_int(), C.c_int()
        _check(lib().disb200_lowres_dims(self.h, C.byref(w), C.byref(h), C.byref(nop)))
        return h.value, w.value, nop.value

    def run_host_lowres(self, A: np.ndarray, B: np.ndarray, out: np.ndarray | None = None) -> np.ndarray:
        """Like run_host, but returns the flow of level sc_l (n, H >> sc_l, W >> sc_l, nop) in level pixel units."""
        n = A.shape[0]
        assert A.dtype == np.uint8 and B.dtype == np.uint8 and A.flags.c_contiguous and B.flags.c_contiguous
        assert A.size == n * self.h0 * self.w0 * self.p.channels and B.shape == A.shape
        if out is None:
            out = np.empty((n,) + self.lowres_dims(), np.float32)
        _check(lib().disb200_run_batch_host_lowres(self.h, _ptr(A, C.c_uint8), _ptr(B, C.c_uint8), n, _ptr(out)))
        return out

    def run_sequence_lowres(self, frames: np.ndarray, out: np.ndarray | None = None) -> np.ndarray:
        nf = frames.shape[0]
        assert frames.dtype == np.uint8 and frames.flags.c_contiguous and frames.size == nf * self.h0 * self.w0 * self.p.channels
        if out is None:
            out = np.empty((nf - 1,) + self.lowres_dims(), np.float32)
        _check(lib().disb200_run_sequence_host_lowres(self.h, _ptr(frames, C.c_uint8), nf, _ptr(out)))
        return out

    def run_host_lowres_ptr(self, a_ptr: int, b_ptr: int, n: int, out_ptr: int):
        _check(lib().disb200_run_batch_host_lowres(self.h, C.c_void_p(a_ptr), C.c_void_p(b_ptr), n, C.c_void_p(out_ptr)))

    def run_sequence_host_lowres_ptr(self, frames_ptr: int, nframes: int, out_ptr: int):
        _check(lib().disb200_run_sequence_host_lowres(self.h, C.c_void_p(frames_ptr), nframes, C.c_void_p(out_ptr)))

    def run_sequence_host_ptr(self, frames_ptr: int, nframes: int, out_ptr: int):
        _check(lib().disb200_run_sequence_host(self.h, C.c_void_p(frames_ptr), nframes, C.c_void_p(out_ptr)))

    def run_sequence_device(self, dFrames: int, nframes: int, dOut: int):
        _check(lib().disb200_run_sequence_device(self.h, C.c_void_p(dFrames), nframes, C.c_void_p(dOut)))

    def run_host_ptr(self, a_ptr: int, b_ptr: int, n: int, out_ptr: int):
        _check(lib().disb200_run_batch_host(self.h, C.c_void_p(a_ptr), C.c_void_p(b_ptr), n, C.c_void_p(out_ptr)))

    def run_list(self, As, Bs):
        n = len(As)
        outs = [np.empty((self.h0, self.w0, self.p.nop), np.float32) for _ in range(n)]
        pa = (C.POINTER(C.c_uint8) * n)(*[_ptr(a, C.c_uint8) for a in As])
        pb = (C.POINTER(C.c_uint8) * n)(*[_ptr(b, C.c_uint8) for b in Bs])
        po = (C.POINTER(C.c_float) * n)(*[_ptr(o) for o in outs])
        _check(lib().disb200_run_batch(self.h, pa, pb, n, po))
        return outs

    # -- device buffers (raw pointers, e.g. torch tensor .data_ptr()) --------------------------
    def run_device(self, dA: int, dB: int, n: int, dOut: int):
        _check(lib().disb200_run_batch_device(self.h, C.c_void_p(dA), C.c_void_p(dB), n, C.c_void_p(dOut)))

    def sync(self):
        _check(lib().disb200_sync(self.h))

    def timer_begin(self):
        _check(lib().disb200_timer_begin(self.h))

    def timer_end(self) -> float:
        ms = C.c_float()
        _check(lib().disb200_timer_end(self.h, C.byref(ms)))
        return ms.value

    def launch_count(self) -> int:
        return lib().disb200_last_launch_count(self.h)

    def set_stage_timing(self, on: bool):
        _check(lib().disb200_set_stage_timing(self.h, int(on)))

    def stage_ms(self):
        out = {}
        for i, name in enumerate(STAGES):
            ms = C.c_float()
            _check(lib().disb200_stage_ms(self.h, i, C.byref(ms)))
            out[name] = ms.value
        return out

    # -- intermediates ---------------------------------------------------------------------
    def scale_ms(self, level: int):
        """(search_ms, densify_ms, refine_ms, npatches) of one pyramid level in the last stage-timed run (oflow.cpp:297)."""
        a, b, c, n = C.c_float(), C.c_float(), C.c_float(), C.c_int()
        _check(lib().disb200_scale_ms(self.h, level, C.byref(a), C.byref(b), C.byref(c), C.byref(n)))
        return a.value, b.value, c.value, n.value

    def level_shape(self, level):
        return self.H >> level, self.W >> level

    def fetch_pyramid(self, which: int, level: int, pair: int = 0):
        hl, wl = self.level_shape(level)
        pad = self.p.patchsz
        outs = []
        for kind in range(3):
            a = np.empty((hl + 2 * pad, wl + 2 * pad, self.p.channels), np.float32)
            _check(lib().disb200_fetch(self.h, kind, which, level, pair, _ptr(a), C.c_size_t(a.nbytes)))
            outs.append(a)
        return outs

    def fetch_flow(self, which: int, level: int, pair: int = 0):
        hl, wl = self.level_shape(level)
        a = np.empty((hl, wl, self.p.nop), np.float32)
        _check(lib().disb200_fetch(self.h, 3, which, level, pair, _ptr(a), C.c_size_t(a.nbytes)))
        return a

    def fetch_patches(self, which: int = 0, pair: int = 0):
        hl, wl = self.level_shape(self.p.sc_l)
        steps = max(1, int(np.floor(np.float32(self.p.patchsz) * (np.float32(1) - np.float32(self.p.patove)))))
        n = int(np.ceil(np.float32(wl) / np.float32(steps))) * int(np.ceil(np.float32(hl) / np.float32(steps)))
        nv = self.p.channels * self.p.patchsz ** 2
        r = {"p_iter": np.empty((n, self.p.nop), np.float32), "pt_iter": np.empty((n, 2), np.float32),
             "pweight": np.empty((n, nv), np.float32)}
        for kind, key in ((4, "p_iter"), (5, "pt_iter"), (6, "pweight")):
            _check(lib().disb200_fetch(self.h, kind, which, self.p.sc_l, pair, _ptr(r[key]), C.c_size_t(r[key].nbytes)))
        return r


class Tiled:
    """disb200_tiled_*: one frame pair split spatially over several GPUs (device ids may repeat: strips on one GPU)."""

    def __init__(self, params: DisParams, w0: int, h0: int, devices):
        self.p = params.copy()
        self.w0, self.h0, self.devices = w0, h0, list(devices)
        self.h = C.c_void_p()
        dv = (C.c_int * len(self.devices))(*self.devices)
        _check(lib().disb200_tiled_create(C.byref(self.p), w0, h0, len(self.devices), dv, C.byref(self.h)))

    def close(self):
        if self.h:
            lib().disb200_tiled_destroy(self.h)
            self.h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def run(self, A: np.ndarray, B: np.ndarray) -> np.ndarray:
        A, B = np.ascontiguousarray(A, np.uint8), np.ascontiguousarray(B, np.uint8)
        out = np.empty((self.h0, self.w0, self.p.nop), np.float32)
        _check(lib().disb200_tiled_run_host(self.h, _ptr(A, C.c_uint8), _ptr(B, C.c_uint8), _ptr(out)))
        return out

    def time_ms(self, reps: int = 5) -> float:
        ms = C.c_float()
        _check(lib().disb200_tiled_time(self.h, reps, C.byref(ms)))
        return ms.value

    def plan(self, level: int):
        """-> (tiled, row boundaries, patch-row boundaries) of one pyramid level"""
        n = len(self.devices)
        t, rows, gy = C.c_int(), (C.c_int * (n + 1))(), (C.c_int * (n + 1))()
        _check(lib().disb200_tiled_plan(self.h, level, C.byref(t), rows, gy))
        return bool(t.value), list(rows), list(gy)

    def launch_count(self) -> int:
        return lib().disb200_tiled_last_launch_count(self.h)


class DisTuning(C.Structure):
    """Mirror of struct DisTuning (schedule selection; 0 = automatic)."""
    _fields_ = [("streams", C.c_int), ("host_chunk", C.c_int), ("tv_schedule", C.c_int), ("sorb_max", C.c_int), ("patch_map", C.c_int),
                ("antiphase", C.c_int)]


TV_AUTO, TV_FUSED, TV_SPLIT, TV_GENERAL, TV_SOLVEF, TV_TEAM = 0, 1, 2, 3, 4, 5


def set_tuning(**kw):
    """disb200_set_tuning; no arguments restores the automatic choices.  Read by handles created afterwards."""
    if not kw:
        _check(lib().disb200_set_tuning(None))
        return
    t = DisTuning(**kw)
    _check(lib().disb200_set_tuning(C.byref(t)))


def _pyr_ptrs(params, A, B):
    fp = C.POINTER(C.c_float)
    ptrs = []
    for d in (*A, *B):
        arr = (fp * (params.sc_f + 1))()
        for l, a in d.items():
            assert a.dtype == np.float32 and a.flags.c_contiguous
            arr[l] = a.ctypes.data_as(fp)
        ptrs.append(arr)
    return ptrs


def ofclass(params: DisParams, A, B, W: int, H: int, initflow=None) -> np.ndarray:
    """OFClass-level entry: A, B are (img, dx, dy) dict triples {level: ndarray} of padded arrays."""
    fp = C.POINTER(C.c_float)
    ptrs = []
    for d in (*A, *B):
        arr = (fp * (params.sc_f + 1))()
        for l, a in d.items():
            assert a.dtype == np.float32 and a.flags.c_contiguous
            arr[l] = a.ctypes.data_as(fp)
        ptrs.append(arr)
    out = np.zeros((H >> params.sc_l, W >> params.sc_l, params.nop), np.float32)
    _check(lib().disb200_ofclass(C.byref(params), *ptrs, W, H, _ptr(out), _ptr(initflow)))
    return out


def aggregate_dense(params: DisParams, w_l: int, h_l: int, p_iter, pweight, cg=None) -> np.ndarray:
    """AggregateFlowDense cut (patchgrid.cpp:213-397); cg = (p_iter, pt_iter, pweight) of the complementary grid."""
    out = np.empty((h_l, w_l, params.nop), np.float32)
    arrs = [np.ascontiguousarray(x, np.float32) for x in (p_iter, pweight) + (tuple(cg) if cg else ())]
    ptrs = [_ptr(x) for x in arrs] + ([None] * 3 if not cg else [])
    _check(lib().disb200_aggregate_dense(C.byref(params), w_l, h_l, *ptrs, _ptr(out), None))
    return out


STATIC_SCRATCH_BYTES = 1 << 20     # DISB200_STATIC_SCRATCH_BYTES


def aggregate_dense_static(params: DisParams, w_l: int, h_l: int, d_p_iter: int, d_pweight: int, d_cg=None, d_flow: int = 0,
                           stream: int = 0):
    """disb200_aggregate_dense_static: the allocation-free variant of aggregate_dense.  Every array argument is a DEVICE
    address (an int, e.g. torch.Tensor.data_ptr()); d_cg = (p_iter, pt_iter, pweight) addresses or None.  Enqueues on
    `stream` (a cudaStream_t handle, 0 = default stream) and returns without synchronising."""
    vp = lambda x: C.c_void_p(x) if x else None
    cg = tuple(d_cg) if d_cg else (0, 0, 0)
    _check(lib().disb200_aggregate_dense_static(C.byref(params), w_l, h_l, vp(d_p_iter), vp(d_pweight), vp(cg[0]), vp(cg[1]), vp(cg[2]),
                                                vp(d_flow), vp(stream)))


def patch_count(params: DisParams, w_l: int, h_l: int) -> int:
    n = C.c_int()
    _check(lib().disb200_aggregate_dense(C.byref(params), w_l, h_l, None, None, None, None, None, None, C.byref(n)))
    return n.value


def varref_level(params: DisParams, im_a, im_b, w_l: int, h_l: int, level: int, camlr: int, flow) -> np.ndarray:
    flow = np.ascontiguousarray(flow, np.float32).copy()
    _check(lib().disb200_varref_level(C.byref(params), _ptr(im_a), _ptr(im_b), w_l, h_l, level, camlr, _ptr(flow)))
    return flow


def save_flo(path: str, flow: np.ndarray):
    flow = np.ascontiguousarray(flow, np.float32)
    h, w = flow.shape[:2]
    nc = 1 if flow.ndim == 2 else flow.shape[2]
    _check(lib().disb200_save_flo(path.encode(), _ptr(flow), w, h, nc))


def read_flo(path: str, nc: int = 2) -> np.ndarray:
    """ReadFlowFile (run_dense.cpp:84-128): (h, w, nc) float32, or (h, w) for nc == 1."""
    data = C.POINTER(C.c_float)()
    w, h = C.c_int(), C.c_int()
    _check(lib().disb200_read_flo(path.encode(), nc, C.byref(data), C.byref(w), C.byref(h)))
    try:
        arr = np.ctypeslib.as_array(data, shape=(h.value * w.value * nc,)).copy()
    finally:
        lib().disb200_free(data)
    return arr.reshape((h.value, w.value) if nc == 1 else (h.value, w.value, nc))


def save_pfm(path: str, disp: np.ndarray):
    disp = np.ascontiguousarray(disp, np.float32)
    h, w = disp.shape[:2]
    _check(lib().disb200_save_pfm(path.encode(), _ptr(disp), w, h))


def read_pnm(path: str, channels: int) -> np.ndarray:
    return _read(lib().disb200_read_pnm, path, channels)


def read_image(path: str, channels: int) -> np.ndarray:
    """cv::imread stand-in (run_dense.cpp:209-210) for PGM / PPM / PNG: uint8 (h, w) grey or (h, w, 3) BGR."""
    return _read(lib().disb200_read_image, path, channels)


class Writer:
    """Pool of host threads writing .flo / .pfm files (disb200_writer_*); submit copies and returns."""

    def __init__(self, threads: int = 4, max_queued_bytes: int = 0):
        self.w = C.c_void_p()
        _check(lib().disb200_writer_create(threads, C.c_size_t(max_queued_bytes), C.byref(self.w)))

    def submit_flo(self, path: str, flow: np.ndarray):
        flow = np.ascontiguousarray(flow, np.float32)
        h, w = flow.shape[:2]
        _check(lib().disb200_writer_submit_flo(self.w, path.encode(), _ptr(flow), w, h, 1 if flow.ndim == 2 else flow.shape[2]))

    def submit_pfm(self, path: str, disp: np.ndarray):
        disp = np.ascontiguousarray(disp, np.float32)
        h, w = disp.shape[:2]
        _check(lib().disb200_writer_submit_pfm(self.w, path.encode(), _ptr(disp), w, h))

    def wait(self):
        _check(lib().disb200_writer_wait(self.w))

    def close(self):
        if self.w:
            lib().disb200_writer_destroy(self.w)
            self.w = C.c_void_p()

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()


def _read(fn, path: str, channels: int) -> np.ndarray:
    data = C.POINTER(C.c_uint8)()
    w, h = C.c_int(), C.c_int()
    _check(fn(path.encode(), channels, C.byref(data), C.byref(w), C.byref(h)))
    n = w.value * h.value * channels
    arr = np.ctypeslib.as_array(data, shape=(n,)).copy()
    lib().disb200_free(data)
    return arr.reshape((h.value, w.value) if channels == 1 else (h.value, w.value, 3))
