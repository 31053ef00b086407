"""Shared test helpers (test infrastructure)."""
import numpy as np

from oracle import bindings as ob


def to_orc(p):
    """DisParams (product) -> OrcParams (oracle)."""
    q = ob.OrcParams()
    for f, _ in ob.OrcParams._fields_:
        setattr(q, f, getattr(p, {"psz": "patchsz", "noc": "channels"}.get(f, f)))
    return q


def to_dis(dis, q, **kw):
    """OrcParams (oracle) -> DisParams (product)."""
    p = dis.DisParams()
    for f, _ in ob.OrcParams._fields_:
        setattr(p, {"psz": "patchsz", "noc": "channels"}.get(f, f), getattr(q, f))
    p.sor_mode = 0
    return p.copy(**kw)


def epe(a, b):
    d = np.sqrt(((np.asarray(a, np.float64) - np.asarray(b, np.float64)) ** 2).sum(-1))
    return float(d.mean()), float(d.max())


def base_orc(mode, noc, **kw):
    p = ob.OrcParams(sc_f=3, sc_l=1, max_iter=12, min_iter=12, dp_thresh=.05, dr_thresh=.95, res_thresh=0., psz=8,
                     patove=.4, usefbcon=0, patnorm=1, costfct=0, usetvref=1, tv_alpha=10., tv_gamma=10., tv_delta=5.,
                     tv_innerit=1, tv_solverit=3, tv_sor=1.6, verbosity=0, mode=mode, noc=noc, gradmag=0)
    return p.copy(**kw)


TINY_CASES = [
    # mode, noc, w0, h0, sc_f, sc_l, overrides: coarsest level 1 x 2 (sor_coupled's point-wise fallback, solver.c:86-89),
    # 1 x 1 and 2 x 1 without refinement, zero solver sweeps, a 40 x 24 frame at its automatic scales
    (1, 1, 32, 64, 5, 0, {}), (1, 1, 30, 50, 5, 1, dict(usefbcon=1)), (2, 1, 32, 64, 5, 0, {}), (1, 3, 32, 128, 5, 2, {}),
    (1, 1, 16, 16, 4, 0, dict(usetvref=0)), (1, 1, 32, 16, 4, 0, dict(usetvref=0)),
    (1, 1, 96, 64, 3, 0, dict(tv_solverit=0)), (2, 1, 96, 64, 3, 0, dict(tv_solverit=0)), (1, 1, 40, 24, 1, 0, {}),
]
