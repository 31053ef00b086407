"""TEST / BASELINE INFRASTRUCTURE ONLY - the reference driver's OpenCV steps through python-cv2.

The reference's pre- and post-processing around the OFClass constructor is OpenCV
(run_dense.cpp:131-179 ConstructImgPyramide, :299-312 dimension padding, :407-415 resize +
crop).  The C++ OpenCV the reference links is not in this image; its Python wheel (cv2 4.13,
same library, SIMD build) is.  bench.py's CPU-baseline legs time THESE calls for the pre/post
share, so that the baseline is not padded by the scalar restatement in dis_oracle.cpp (which
stays the parity oracle: it matches cv2 bit for bit on u8-sourced frames, tests/test_oracle_cpu.py).

Only bench.py's cpu_baseline / --impl reference legs and tests/ import this module.
"""
from __future__ import annotations

import numpy as np

try:
    import cv2
    cv2.setNumThreads(1)          # one pair per host worker; OpenCV's own pool would oversubscribe
    HAVE_CV2 = True
except Exception:                 # pragma: no cover - the wheel is part of the image
    cv2 = None
    HAVE_CV2 = False


def pad_frame(img_u8: np.ndarray, sc_f: int):
    """run_dense.cpp:299-312 -> (padded frame, padl, padt)."""
    h0, w0 = img_u8.shape[:2]
    scfct = 1 << sc_f
    padw = (scfct - w0 % scfct) % scfct
    padh = (scfct - h0 % scfct) % scfct
    if padw or padh:
        img_u8 = cv2.copyMakeBorder(img_u8, padh // 2, padh - padh // 2, padw // 2, padw - padw // 2, cv2.BORDER_REPLICATE)
    return img_u8, padw // 2, padh // 2


def pyramid(img_u8: np.ndarray, sc_f: int, pad: int, gradmag: bool = False):
    """ConstructImgPyramide (run_dense.cpp:131-179) of one already dimension-padded frame: ALL levels
    0..sc_f as the reference builds them (it never reads the ones below sc_l, oflow.cpp:182, but it
    pays for them).  -> ({level: img}, {level: dx}, {level: dy}), padded by `pad`, float32."""
    f = img_u8.astype(np.float32)                                           # convertTo CV_32F, :327-328
    I, X, Y = {}, {}, {}
    prev = None
    for l in range(sc_f + 1):
        if l == 0:
            if gradmag:                                                     # SELECTCHANNEL == 2, :139-148
                dx = cv2.Sobel(f, cv2.CV_32F, 1, 0, ksize=3, scale=1 / 8.0, delta=0, borderType=cv2.BORDER_DEFAULT)
                dy = cv2.Sobel(f, cv2.CV_32F, 0, 1, ksize=3, scale=1 / 8.0, delta=0, borderType=cv2.BORDER_DEFAULT)
                cur = cv2.sqrt(dx * dx + dy * dy)
            else:
                cur = f.copy()                                              # clone(), :137
        else:
            cur = cv2.resize(prev, None, fx=.5, fy=.5, interpolation=cv2.INTER_LINEAR)   # :151
        prev = cur
        dx = cv2.Sobel(cur, cv2.CV_32F, 1, 0, ksize=3, scale=1 / 8.0, delta=0, borderType=cv2.BORDER_DEFAULT)   # :157
        dy = cv2.Sobel(cur, cv2.CV_32F, 0, 1, ksize=3, scale=1 / 8.0, delta=0, borderType=cv2.BORDER_DEFAULT)   # :158
        I[l] = cv2.copyMakeBorder(cur, pad, pad, pad, pad, cv2.BORDER_REPLICATE)                               # :167
        X[l] = cv2.copyMakeBorder(dx, pad, pad, pad, pad, cv2.BORDER_CONSTANT, value=0)                        # :172
        Y[l] = cv2.copyMakeBorder(dy, pad, pad, pad, pad, cv2.BORDER_CONSTANT, value=0)                        # :173
    return I, X, Y


def upsample_crop(low: np.ndarray, sc_l: int, padl: int, padt: int, w0: int, h0: int) -> np.ndarray:
    """run_dense.cpp:407-415: flow *= 2^sc_l, cv::resize INTER_LINEAR by 2^sc_l, crop the padding."""
    nop = low.shape[2]
    out = low
    if sc_l != 0:
        s = float(1 << sc_l)
        out = cv2.resize(low * np.float32(s), None, fx=s, fy=s, interpolation=cv2.INTER_LINEAR)
        if out.ndim == 2:
            out = out[:, :, None]
    return np.ascontiguousarray(out[padt:padt + h0, padl:padl + w0, :nop])
