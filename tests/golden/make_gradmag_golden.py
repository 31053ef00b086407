"""Golden fixtures for the gradient-magnitude input mode (SELECTCHANNEL == 2, run_dense.cpp:139-148).

  python tests/golden/make_gradmag_golden.py      (build container only: needs cv2 and /root/reference)

gradmag.npz holds
* the OpenCV steps of ConstructImgPyramide for that mode, evaluated with python-cv2 4.13 on a seeded u8
  frame: Sobel/8 of the padded float frame, dx.mul(dx) + dy.mul(dy), cv::sqrt, then the .5 resize cascade
  and the Sobel gradients of each level.  Unlike the intensity pyramid the cascade now runs on arbitrary
  floats, where the summation order shows: OpenCV's own code computes ((a+b)+(c+d))*0.25
  (`*_l1..3`, cv2.setUseOptimized(False)); wheels that dispatch cv::resize to IPP compute
  lerp(lerp(a,b), lerp(c,d)) and differ in the last bit of ~8 % of the pixels (`*_ipp_l1..3`, kept to
  document the spread: OpenCV is an unpinned dependency of the reference).  The parity target is
  OpenCV's own code;
* the low-resolution flow of the reference's own sources compiled with -DSELECTCHANNEL=2
  (oracle/_ref/libref_{OF,DE}_GRAD.so) on a seeded pair.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)


def cv2_gradmag(u8, padt, padb, padl, padr, levels, optimized):
    import cv2
    cv2.setUseOptimized(optimized)
    f = cv2.copyMakeBorder(u8, padt, padb, padl, padr, cv2.BORDER_REPLICATE).astype(np.float32)
    dx = cv2.Sobel(f, cv2.CV_32F, 1, 0, ksize=3, scale=1 / 8.0, delta=0, borderType=cv2.BORDER_DEFAULT)
    dy = cv2.Sobel(f, cv2.CV_32F, 0, 1, ksize=3, scale=1 / 8.0, delta=0, borderType=cv2.BORDER_DEFAULT)
    mag = cv2.sqrt(cv2.add(cv2.multiply(dx, dx), cv2.multiply(dy, dy)))
    out = {"l0": mag}
    lv = mag
    for l in range(1, levels + 1):
        lv = cv2.resize(lv, None, fx=.5, fy=.5, interpolation=cv2.INTER_LINEAR)
        out["l%d" % l] = lv
        out["dx%d" % l] = cv2.Sobel(lv, cv2.CV_32F, 1, 0, ksize=3, scale=1 / 8.0, delta=0, borderType=cv2.BORDER_DEFAULT)
        out["dy%d" % l] = cv2.Sobel(lv, cv2.CV_32F, 0, 1, ksize=3, scale=1 / 8.0, delta=0, borderType=cv2.BORDER_DEFAULT)
    cv2.setUseOptimized(True)
    return out


def main():
    import cv2
    from oracle import bindings as ob
    from of_dis_b200.synth import make_pair
    rng = np.random.default_rng(17)
    out = {"cv2_version": np.array(cv2.__version__)}
    u8 = rng.integers(0, 256, (44, 72), dtype=np.uint8)
    out["g_u8"] = u8
    for k, v in cv2_gradmag(u8, 2, 2, 4, 4, 3, False).items():
        out["g_" + k] = v
    ipp = cv2_gradmag(u8, 2, 2, 4, 4, 3, True)
    for l in (1, 2, 3):
        out["g_ipp_l%d" % l] = ipp["l%d" % l]
    assert np.array_equal(ipp["l0"], out["g_l0"])

    ob.build()
    port = ob.Port()
    for mode, name in ((1, "of_grad_op2like"), (2, "de_grad_op2like")):
        A, B, gt = make_pair(90, 150, channels=1, mode="OF" if mode == 1 else "DE", seed=11)
        p = ob.OrcParams(sc_f=3, sc_l=1, max_iter=12, min_iter=12, dp_thresh=.05, dr_thresh=.95, res_thresh=0., psz=8,
                         patove=.4, usefbcon=0, patnorm=1, costfct=0, usetvref=1, tv_alpha=10., tv_gamma=10.,
                         tv_delta=5., tv_innerit=1, tv_solverit=3, tv_sor=1.6, verbosity=0, mode=mode, noc=1, gradmag=1)
        W, H, padl, padt = port.pad_dims(150, 90, p.sc_f)
        PA = port.pyramid(A, p.sc_f, p.sc_l, p.psz, gradmag=True)
        PB = port.pyramid(B, p.sc_f, p.sc_l, p.psz, gradmag=True)
        low = ob.Ref(mode, 1, gradmag=True).ofclass(p, PA, PB, W, H)
        out[name + "_A"], out[name + "_B"], out[name + "_gt"], out[name + "_low"] = A, B, gt, low
        print(name, low.shape, float(np.abs(low).mean()))
    np.savez_compressed(os.path.join(HERE, "gradmag.npz"), **out)
    print("gradmag.npz:", len(out), "arrays")


if __name__ == "__main__":
    main()
