"""Generates the committed golden fixtures (run once in the build container, not on the GPU box).

  python tests/golden/make_golden.py

* cv2_steps.npz   - the OpenCV calls the reference delegates to (run_dense.cpp:131-179,299-312,
                    407-415), evaluated with python-cv2 (4.13 here) on small seeded inputs.
* ref_flows.npz   - outputs of the reference's OWN code (oracle/_ref, compiled unmodified from
                    /root/reference) on small seeded pairs, for the four binaries and a few
                    parameter variants, with the u8 inputs stored alongside.
Neither cv2 nor /root/reference is needed to run the tests: only these files are read.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)


def cv2_steps():
    import cv2
    rng = np.random.default_rng(7)
    out = {"cv2_version": np.array(cv2.__version__)}
    for tag, shape in (("g", (44, 72)), ("c", (44, 72, 3))):
        u8 = rng.integers(0, 256, shape, dtype=np.uint8)
        # run_dense.cpp:309 copyMakeBorder(REPLICATE) to a multiple of 2^sc_f = 16 -> 48 x 80
        padded = cv2.copyMakeBorder(u8, 2, 2, 4, 4, cv2.BORDER_REPLICATE)
        f = padded.astype(np.float32)
        out[tag + "_u8"] = u8
        out[tag + "_l0"] = f
        lv = f
        for l in (1, 2, 3):
            lv = cv2.resize(lv, None, fx=.5, fy=.5, interpolation=cv2.INTER_LINEAR)
            out["%s_l%d" % (tag, l)] = lv
            out["%s_dx%d" % (tag, l)] = cv2.Sobel(lv, cv2.CV_32F, 1, 0, ksize=3, scale=1 / 8.0, delta=0, borderType=cv2.BORDER_DEFAULT)
            out["%s_dy%d" % (tag, l)] = cv2.Sobel(lv, cv2.CV_32F, 0, 1, ksize=3, scale=1 / 8.0, delta=0, borderType=cv2.BORDER_DEFAULT)
        out[tag + "_pad2_rep"] = cv2.copyMakeBorder(out[tag + "_l2"], 8, 8, 8, 8, cv2.BORDER_REPLICATE)
        out[tag + "_pad2_zero"] = cv2.copyMakeBorder(out[tag + "_dx2"], 8, 8, 8, 8, cv2.BORDER_CONSTANT, value=0)
    for nop in (1, 2):
        for sc_l in (1, 2, 3):
            fl = (rng.standard_normal((12, 20, nop)) * 3).astype(np.float32)
            k = 1 << sc_l
            up = cv2.resize(fl * np.float32(k), None, fx=k, fy=k, interpolation=cv2.INTER_LINEAR)
            if up.ndim == 2:
                up = up[..., None]
            out["up_in_%d_%d" % (nop, sc_l)] = fl
            # crop Rect(padl, padt, w0, h0) with padl = 1, padt = 2 (run_dense.cpp:415)
            out["up_out_%d_%d" % (nop, sc_l)] = up[2:up.shape[0] - 3, 1:up.shape[1] - 2]
    np.savez_compressed(os.path.join(HERE, "cv2_steps.npz"), **out)
    print("cv2_steps.npz:", len(out), "arrays")


REF_CASES = {
    # name: (mode, noc, h0, w0, overrides)  -- explicit 20-parameter style
    "of_int_op2like": (1, 1, 90, 150, dict(sc_f=3, sc_l=1, max_iter=12, min_iter=12, psz=8, patove=.4, usetvref=1)),
    "of_int_notv": (1, 1, 90, 150, dict(sc_f=3, sc_l=1, max_iter=16, min_iter=16, psz=8, patove=.3, usetvref=0)),
    "of_int_psz12": (1, 1, 90, 150, dict(sc_f=2, sc_l=0, max_iter=16, min_iter=16, psz=12, patove=.75, usetvref=1)),
    "of_int_l1_fb": (1, 1, 90, 150, dict(sc_f=3, sc_l=1, max_iter=12, min_iter=12, psz=8, patove=.4, usetvref=1, costfct=1, usefbcon=1)),
    "of_int_huber_early": (1, 1, 90, 150, dict(sc_f=3, sc_l=1, max_iter=24, min_iter=2, psz=8, patove=.4, usetvref=1, costfct=2, res_thresh=.5)),
    "of_rgb_op2like": (1, 3, 90, 150, dict(sc_f=3, sc_l=1, max_iter=12, min_iter=12, psz=8, patove=.4, usetvref=1)),
    "de_int_op2like": (2, 1, 90, 150, dict(sc_f=3, sc_l=1, max_iter=12, min_iter=12, psz=8, patove=.4, usetvref=1)),
    "de_rgb_fb": (2, 3, 90, 150, dict(sc_f=3, sc_l=1, max_iter=12, min_iter=12, psz=8, patove=.4, usetvref=1, usefbcon=1)),
}


def base_params(ob, mode, noc, **kw):
    p = ob.OrcParams(sc_f=3, sc_l=1, max_iter=12, min_iter=12, dp_thresh=.05, dr_thresh=.95, res_thresh=0., psz=8,
                     patove=.4, usefbcon=0, patnorm=1, costfct=0, usetvref=1, tv_alpha=10., tv_gamma=10., tv_delta=5.,
                     tv_innerit=1, tv_solverit=3, tv_sor=1.6, verbosity=0, mode=mode, noc=noc)
    return p.copy(**kw)


def ref_flows():
    from oracle import bindings as ob
    from of_dis_b200.synth import make_pair
    ob.build()
    port = ob.Port()
    out = {}
    for name, (mode, noc, h0, w0, kw) in REF_CASES.items():
        A, B, gt = make_pair(h0, w0, channels=noc, mode="OF" if mode == 1 else "DE", seed=11)
        p = base_params(ob, mode, noc, **kw)
        W, H, padl, padt = port.pad_dims(w0, h0, p.sc_f)
        PA = port.pyramid(A, p.sc_f, p.sc_l, p.psz)
        PB = port.pyramid(B, p.sc_f, p.sc_l, p.psz)
        ref = ob.Ref(mode, noc)
        low = ref.ofclass(p, PA, PB, W, H)
        out[name + "_A"], out[name + "_B"], out[name + "_gt"], out[name + "_low"] = A, B, gt, low
        print(name, low.shape, float(np.abs(low).mean()))
    np.savez_compressed(os.path.join(HERE, "ref_flows.npz"), **out)


if __name__ == "__main__":
    cv2_steps()
    ref_flows()
