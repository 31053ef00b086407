"""TEST INFRASTRUCTURE ONLY - ctypes bindings to the parity oracles.

* ``Port``  -> oracle/libdisoracle.so  (CPU restatement, oracle/dis_oracle.cpp)
* ``Ref``   -> oracle/_ref/libref_{OF,DE}_{INT,RGB,GRAD}.so (the reference's own sources compiled
               unmodified by oracle/Makefile; exists only where /root/reference was present at
               build time or the prebuilt .so travelled with the snapshot)

Only tests/, __graft_entry__.smoke() and bench.py's CPU-baseline legs may import this module.
The product package (of_dis_b200) never does.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REF_SRC = "/root/reference"

f32p = C.POINTER(C.c_float)
i32p = C.POINTER(C.c_int)
u8p = C.POINTER(C.c_uint8)


class OrcParams(C.Structure):
    _fields_ = [
        ("sc_f", C.c_int), ("sc_l", C.c_int), ("max_iter", C.c_int), ("min_iter", C.c_int),
        ("dp_thresh", C.c_float), ("dr_thresh", C.c_float), ("res_thresh", C.c_float),
        ("psz", C.c_int), ("patove", C.c_float),
        ("usefbcon", C.c_int), ("patnorm", C.c_int), ("costfct", C.c_int), ("usetvref", C.c_int),
        ("tv_alpha", C.c_float), ("tv_gamma", C.c_float), ("tv_delta", C.c_float),
        ("tv_innerit", C.c_int), ("tv_solverit", C.c_int), ("tv_sor", C.c_float),
        ("verbosity", C.c_int), ("mode", C.c_int), ("noc", C.c_int), ("gradmag", C.c_int),
    ]

    def copy(self, **kw):
        q = OrcParams()
        C.memmove(C.byref(q), C.byref(self), C.sizeof(OrcParams))
        for k, v in kw.items():
            setattr(q, k, v)
        return q


def _p(a, t=f32p):
    return None if a is None else a.ctypes.data_as(t)


def build(port: bool = True, ref: bool = True) -> None:
    """Compile the checkers (building the checker is not using it)."""
    if port:
        subprocess.check_call(["make", "-C", HERE, "port"], stdout=subprocess.DEVNULL)
    if ref and os.path.isdir(REF_SRC):
        subprocess.check_call(["make", "-C", HERE, "ref"], stdout=subprocess.DEVNULL)


def variant_name(mode: int, noc: int, gradmag: bool = False) -> str:
    return ("OF" if mode == 1 else "DE") + "_" + ("GRAD" if gradmag else "INT" if noc == 1 else "RGB")


def have_ref(mode: int = 1, noc: int = 1, gradmag: bool = False) -> bool:
    return os.path.exists(os.path.join(HERE, "_ref", "libref_%s.so" % variant_name(mode, noc, gradmag)))


def pyr_ptrs(levels: dict, sc_f: int):
    """levels: {level: ndarray}; returns a (sc_f+1)-array of float* (null for absent levels)."""
    arr = (f32p * (sc_f + 1))()
    for l, a in levels.items():
        arr[l] = a.ctypes.data_as(f32p)
    return arr


class Port:
    """CPU restatement (oracle/dis_oracle.cpp)."""

    def __init__(self):
        path = os.path.join(HERE, "libdisoracle.so")
        if not os.path.exists(path):
            build(port=True, ref=False)
        self.lib = C.CDLL(path)
        self.lib.orc_run_dense_timed.restype = None

    def operating_point(self, op: int, width: int, mode: int = 1, noc: int = 1) -> OrcParams:
        p = OrcParams()
        self.lib.orc_operating_point(op, width, mode, noc, C.byref(p))
        return p

    def set_sor_order(self, order: int):
        """0 = raster order (reference); 1 = red-black (restatement of the product's DISB200_SOR_REDBLACK)."""
        self.lib.orc_set_sor_order(int(order))

    def pad_dims(self, w0, h0, sc_f):
        W, H, pl, pt = C.c_int(), C.c_int(), C.c_int(), C.c_int()
        self.lib.orc_pad_dims(w0, h0, sc_f, C.byref(W), C.byref(H), C.byref(pl), C.byref(pt))
        return W.value, H.value, pl.value, pt.value

    def gradient_magnitude(self, src):
        src = np.ascontiguousarray(src, np.float32)
        dst = np.empty_like(src)
        self.lib.orc_gradient_magnitude(_p(src), src.shape[1], src.shape[0], _p(dst))
        return dst

    def pyramid_level(self, img_u8, sc_f, level, pad, gradmag=False):
        h0, w0 = img_u8.shape[:2]
        Cn = 1 if img_u8.ndim == 2 else img_u8.shape[2]
        W, H, _, _ = self.pad_dims(w0, h0, sc_f)
        wl, hl = W >> level, H >> level
        shp = (hl + 2 * pad, wl + 2 * pad, Cn)
        out = [np.empty(shp, np.float32) for _ in range(3)]
        img_u8 = np.ascontiguousarray(img_u8)
        self.lib.orc_set_pyramid_gradmag(int(bool(gradmag)))
        try:
            self.lib.orc_pyramid_level(_p(img_u8, u8p), w0, h0, Cn, sc_f, level, pad, *[_p(o) for o in out])
        finally:
            self.lib.orc_set_pyramid_gradmag(0)
        return out

    def pyramid(self, img_u8, sc_f, sc_l, pad, gradmag=False):
        """-> ({level: img}, {level: dx}, {level: dy}); one cascade per frame"""
        h0, w0 = img_u8.shape[:2]
        Cn = 1 if img_u8.ndim == 2 else img_u8.shape[2]
        W, H, _, _ = self.pad_dims(w0, h0, sc_f)
        I, X, Y = {}, {}, {}
        for l in range(sc_l, sc_f + 1):
            shp = ((H >> l) + 2 * pad, (W >> l) + 2 * pad, Cn)
            I[l], X[l], Y[l] = (np.empty(shp, np.float32) for _ in range(3))
        img_u8 = np.ascontiguousarray(img_u8)
        self.lib.orc_set_pyramid_gradmag(int(bool(gradmag)))
        try:
            self.lib.orc_pyramid_all(_p(img_u8, u8p), w0, h0, Cn, sc_f, sc_l, pad, pyr_ptrs(I, sc_f), pyr_ptrs(X, sc_f), pyr_ptrs(Y, sc_f))
        finally:
            self.lib.orc_set_pyramid_gradmag(0)
        return I, X, Y

    def halve(self, src):
        src = np.ascontiguousarray(src, np.float32)
        h, w = src.shape[:2]
        Cn = 1 if src.ndim == 2 else src.shape[2]
        dst = np.empty((h // 2, w // 2) + src.shape[2:], np.float32)
        self.lib.orc_halve(_p(src), w, h, Cn, _p(dst))
        return dst

    def sobel(self, src):
        src = np.ascontiguousarray(src, np.float32)
        h, w = src.shape[:2]
        Cn = 1 if src.ndim == 2 else src.shape[2]
        dx, dy = np.empty_like(src), np.empty_like(src)
        self.lib.orc_sobel(_p(src), w, h, Cn, _p(dx), _p(dy))
        return dx, dy

    def upsample_crop(self, flow, sc_l, padl, padt, w0, h0):
        flow = np.ascontiguousarray(flow, np.float32)
        hl, wl, nop = flow.shape
        out = np.empty((h0, w0, nop), np.float32)
        self.lib.orc_upsample_crop(_p(flow), wl, hl, nop, sc_l, padl, padt, w0, h0, _p(out))
        return out

    def ofclass(self, p: OrcParams, A, B, W, H, initflow=None):
        """A, B: (img, dx, dy) dict triples as returned by pyramid()."""
        nop = 2 if p.mode == 1 else 1
        out = np.zeros((H >> p.sc_l, W >> p.sc_l, nop), np.float32)
        ptrs = [pyr_ptrs(d, p.sc_f) for d in (*A, *B)]
        self.lib.orc_ofclass(C.byref(p), *ptrs, W, H, _p(out), _p(initflow))
        return out

    def patchgrid_level(self, p: OrcParams, A, B, w_l, h_l, level, prev_fw=None, prev_bw=None):
        return _patchgrid(self.lib.orc_patchgrid_level, True, p, A, B, w_l, h_l, level, prev_fw, prev_bw)

    def aggregate_dense(self, p: OrcParams, w_l, h_l, p_iter, pweight, cg=None):
        """AggregateFlowDense cut; cg = (p_iter, pt_iter, pweight) of the complementary grid or None."""
        nop = 2 if p.mode == 1 else 1
        out = np.zeros((h_l, w_l, nop), np.float32)
        arrs = [np.ascontiguousarray(x, np.float32) for x in (p_iter, pweight) + (tuple(cg) if cg else ())]
        ptrs = [_p(x) for x in arrs] + ([None] * 3 if not cg else [])
        self.lib.orc_aggregate_dense(C.byref(p), w_l, h_l, *ptrs, _p(out))
        return out

    def varref_level(self, p: OrcParams, a, b, w_l, h_l, level, camlr, flow):
        flow = np.ascontiguousarray(flow, np.float32).copy()
        self.lib.orc_varref_level(C.byref(p), _p(a), _p(b), w_l, h_l, level, camlr, _p(flow))
        return flow

    def run_dense(self, p: OrcParams, A_u8, B_u8, want_lowres=False):
        h0, w0 = A_u8.shape[:2]
        nop = 2 if p.mode == 1 else 1
        out = np.empty((h0, w0, nop), np.float32)
        low = None
        if want_lowres:
            W, H, _, _ = self.pad_dims(w0, h0, p.sc_f)
            low = np.empty((H >> p.sc_l, W >> p.sc_l, nop), np.float32)
        A_u8, B_u8 = np.ascontiguousarray(A_u8), np.ascontiguousarray(B_u8)
        self.lib.orc_run_dense(C.byref(p), _p(A_u8, u8p), _p(B_u8, u8p), w0, h0, _p(out), _p(low))
        return (out, low) if want_lowres else out

    def run_dense_timed(self, p: OrcParams, A_u8, B_u8):
        h0, w0 = A_u8.shape[:2]
        nop = 2 if p.mode == 1 else 1
        out = np.empty((h0, w0, nop), np.float32)
        tc, tp = C.c_double(), C.c_double()
        self.lib.orc_run_dense_timed(C.byref(p), _p(A_u8, u8p), _p(B_u8, u8p), w0, h0, _p(out), C.byref(tc), C.byref(tp))
        return out, tc.value, tp.value


def _patchgrid(fn, is_port, p, A, B, w_l, h_l, level, prev_fw, prev_bw):
    nop = 2 if p.mode == 1 else 1
    steps = max(1, int(np.floor(np.float32(p.psz) * (np.float32(1) - np.float32(p.patove)))))
    nopw = int(np.ceil(np.float32(w_l) / np.float32(steps)))
    noph = int(np.ceil(np.float32(h_l) / np.float32(steps)))
    n = nopw * noph
    nv = p.noc * p.psz * p.psz
    r = {}
    for tag in ("fw", "bw"):
        r["p_iter_" + tag] = np.zeros((n, nop), np.float32)
        r["pt_iter_" + tag] = np.zeros((n, 2), np.float32)
        r["pweight_" + tag] = np.zeros((n, nv), np.float32)
        r["dense_" + tag] = np.zeros((h_l, w_l, nop), np.float32)
    r["cnt_fw"] = np.zeros(n, np.int32)
    fb = bool(p.usefbcon)
    imgs = [_p(x) for x in (*A, *B)]
    tail = [_p(prev_fw), _p(prev_bw), _p(r["p_iter_fw"]), _p(r["pt_iter_fw"]), _p(r["pweight_fw"]),
            _p(r["cnt_fw"], i32p), _p(r["dense_fw"]),
            _p(r["p_iter_bw"]) if fb else None, _p(r["pt_iter_bw"]) if fb else None,
            _p(r["pweight_bw"]) if fb else None, _p(r["dense_bw"]) if fb else None]
    if is_port:
        got = fn(C.byref(p), *imgs, w_l, h_l, level, *tail)
    else:
        got = fn(*imgs, w_l, h_l, level, p.max_iter, p.min_iter, C.c_float(p.dp_thresh), C.c_float(p.dr_thresh),
                 C.c_float(p.res_thresh), p.psz, C.c_float(p.patove), p.usefbcon, p.costfct, p.noc, p.patnorm, *tail)
    assert got == n, (got, n)
    r["nopw"], r["noph"], r["steps"] = nopw, noph, steps
    return r


class Ref:
    """The reference's own code (oracle/_ref), one variant per (SELECTMODE, SELECTCHANNEL); gradmag selects the
    SELECTCHANNEL == 2 build (single band like 1; only the driver's pyramid differs, run_dense.cpp:139-148)."""

    def __init__(self, mode: int = 1, noc: int = 1, gradmag: bool = False):
        path = os.path.join(HERE, "_ref", "libref_%s.so" % variant_name(mode, noc, gradmag))
        if not os.path.exists(path):
            raise FileNotFoundError(path + " (run `make -C oracle ref` where /root/reference exists)")
        self.lib = C.CDLL(path)
        assert self.lib.ref_mode() == mode and self.lib.ref_channels() == (2 if gradmag else noc)
        self.mode, self.noc = mode, noc

    def ofclass(self, p: OrcParams, A, B, W, H, initflow=None, verbosity=0):
        nop = 2 if p.mode == 1 else 1
        out = np.zeros((H >> p.sc_l, W >> p.sc_l, nop), np.float32)
        ptrs = [pyr_ptrs(d, p.sc_f) for d in (*A, *B)]
        F = C.c_float
        self.lib.ref_ofclass(*ptrs, p.psz, _p(out), _p(initflow), W, H, p.sc_f, p.sc_l, p.max_iter, p.min_iter,
                             F(p.dp_thresh), F(p.dr_thresh), F(p.res_thresh), p.psz, F(p.patove), p.usefbcon,
                             p.costfct, p.noc, p.patnorm, p.usetvref, F(p.tv_alpha), F(p.tv_gamma), F(p.tv_delta),
                             p.tv_innerit, p.tv_solverit, F(p.tv_sor), verbosity)
        return out

    def patchgrid_level(self, p: OrcParams, A, B, w_l, h_l, level, prev_fw=None, prev_bw=None):
        return _patchgrid(self.lib.ref_patchgrid_level, False, p, A, B, w_l, h_l, level, prev_fw, prev_bw)

    def varref_level(self, p: OrcParams, a, b, w_l, h_l, level, camlr, flow):
        flow = np.ascontiguousarray(flow, np.float32).copy()
        F = C.c_float
        self.lib.ref_varref_level(_p(a), _p(b), w_l, h_l, p.psz, level, camlr, p.noc, F(p.tv_alpha), F(p.tv_gamma),
                                  F(p.tv_delta), p.tv_innerit, p.tv_solverit, F(p.tv_sor), _p(flow))
        return flow
