"""GPU suite (-m gpu): the CUDA path, called through the C-ABI, against the CPU oracle.

Tolerances (north_star): mean EPE(GPU, CPU) <= 0.01 px for every operating point; max-pixel
bound 0.01 px at op1/op2-like settings and 1.0 px at op3/op4-like settings (rare outlier-test
flips).  Stages whose arithmetic order is reproduced exactly are held to
bit-equality: pyramid, variational refinement (lexicographic SOR), upsample.
"""
import importlib.util
import os

import numpy as np
import pytest

from oracle import bindings as ob
from of_dis_b200.synth import make_pair
from tests.helpers import TINY_CASES, base_orc, epe, to_dis, to_orc

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

MEAN_TOL = 0.01
MAX_TOL_FINE = 0.01   # op1 / op2
MAX_TOL_DENSE = 1.0   # op3 / op4 (and other high-overlap settings)


def _ref_cases():
    spec = importlib.util.spec_from_file_location("make_golden", os.path.join(ROOT, "tests", "golden", "make_golden.py"))
    m = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(m)
    return m.REF_CASES


REF_CASES = _ref_cases()


# ------------------------------------------------------------------------------ stage 1
@pytest.mark.parametrize("ch", [1, 3])
@pytest.mark.parametrize("w0,h0,sc_f,sc_l", [(150, 90, 3, 1), (1024, 436, 5, 3), (96, 64, 2, 0), (333, 217, 4, 2)])
def test_pyramid_bit_exact(gpu, port, ch, w0, h0, sc_f, sc_l):
    A, B, _ = make_pair(h0, w0, channels=ch, seed=4)
    p = to_dis(gpu, base_orc(1, ch, sc_f=sc_f, sc_l=sc_l))
    with gpu.Handle(p, w0, h0, max_batch=2) as hd:
        hd.run_host(np.stack([A, B]), np.stack([B, A]))
        for l in range(sc_l, sc_f + 1):
            want_a, want_b = port.pyramid_level(A, sc_f, l, 8), port.pyramid_level(B, sc_f, l, 8)
            for which, pair, want in ((0, 0, want_a), (1, 0, want_b), (0, 1, want_b), (1, 1, want_a)):
                got = hd.fetch_pyramid(which, l, pair)
                for k in range(3):
                    assert np.array_equal(got[k], want[k]), (l, which, pair, k)


@pytest.mark.parametrize("w0,h0,sc_f,sc_l", [(150, 90, 3, 1), (1024, 436, 5, 3), (96, 64, 2, 0), (333, 217, 4, 2), (150, 90, 2, 2),
                                            (100, 60, 3, 3), (1280, 720, 6, 4)])
def test_pyramid_gradmag_bit_exact(gpu, port, w0, h0, sc_f, sc_l):
    """SELECTCHANNEL == 2 (run_dense.cpp:139-148): the cascade runs on arbitrary floats, so OpenCV's summation order
    shows (pinned by tests/golden/gradmag.npz); sizes cover tail columns (w % 4 != 0) inside the fused cascade
    (sc_l >= sc_f - 1) and in the level-to-level halving."""
    A, B, _ = make_pair(h0, w0, channels=1, seed=4)
    p = to_dis(gpu, base_orc(1, 1, sc_f=sc_f, sc_l=sc_l, gradmag=1))
    with gpu.Handle(p, w0, h0, max_batch=2) as hd:
        hd.run_host(np.stack([A, B]), np.stack([B, A]))
        for l in range(sc_l, sc_f + 1):
            want_a, want_b = port.pyramid_level(A, sc_f, l, 8, gradmag=True), port.pyramid_level(B, sc_f, l, 8, gradmag=True)
            for which, pair, want in ((0, 0, want_a), (1, 0, want_b), (0, 1, want_b), (1, 1, want_a)):
                got = hd.fetch_pyramid(which, l, pair)
                for k in range(3):
                    assert np.array_equal(got[k], want[k]), (l, which, pair, k)
        # the video front end builds each frame's gradient-magnitude pyramid once
        seq = hd.run_sequence(np.stack([A, B, A]))
        assert np.array_equal(seq[0], hd.run_host(A[None], B[None])[0]) and np.array_equal(seq[1], hd.run_host(B[None], A[None])[0])


@pytest.mark.parametrize("mode,name", [(1, "of_grad_op2like"), (2, "de_grad_op2like")])
def test_gradmag_against_reference_golden(gpu, gradmag_golden, mode, name):
    """u8 frames in, against the flow of the reference's own sources built with -DSELECTCHANNEL=2."""
    g = gradmag_golden
    A, B = g[name + "_A"], g[name + "_B"]
    p = to_dis(gpu, base_orc(mode, 1, gradmag=1))
    with gpu.Handle(p, 150, 90, max_batch=1) as hd:
        hd.run_host(A[None], B[None])
        low = hd.fetch_flow(0, p.sc_l, pair=0)
    m, mx = epe(low, g[name + "_low"])
    assert m <= MEAN_TOL and mx <= 0.05, (m, mx)


@pytest.mark.parametrize("mode,w0,h0", [(1, 1024, 436), (2, 1280, 720)])
def test_gradmag_run_dense_parity(gpu, port, mode, w0, h0):
    A, B, gt = make_pair(h0, w0, channels=1, mode="OF" if mode == 1 else "DE", seed=0)
    p = gpu.operating_point(2, w0, mode, 1).copy(verbosity=0, gradmag=1)
    want = port.run_dense(to_orc(p), A, B)
    with gpu.Handle(p, w0, h0, max_batch=2) as hd:
        out = hd.run_host(np.stack([A, B]), np.stack([B, A]))
    m, mx = epe(out[0], want)
    assert m <= MEAN_TOL and mx <= MAX_TOL_FINE, "EPE(GPU,CPU) mean %g max %g" % (m, mx)
    assert epe(out[1], port.run_dense(to_orc(p), B, A))[0] <= MEAN_TOL
    assert abs(epe(out[0], gt)[0] - epe(want, gt)[0]) < 5e-3
    assert not np.array_equal(out[0], port.run_dense(to_orc(p.copy(gradmag=0)), A, B))


# ------------------------------------------------------------------------------ stage 4 alone
@pytest.mark.parametrize("mode,ch", [(1, 1), (1, 3), (2, 1), (2, 3)])
@pytest.mark.parametrize("w0,h0,level", [(1024, 436, 3), (500, 218, 1), (130, 70, 0)])
def test_varref_level_bit_exact(gpu, port, mode, ch, w0, h0, level):
    """Covers widths that are not multiples of 4 (FDF stride padding) and both cameras."""
    A, B, _ = make_pair(h0, w0, channels=ch, mode="OF" if mode == 1 else "DE", seed=1)
    q = base_orc(mode, ch, sc_f=level, sc_l=level)
    p = to_dis(gpu, q)
    W, H, _, _ = port.pad_dims(w0, h0, level)
    a, b = port.pyramid_level(A, level, level, 8)[0], port.pyramid_level(B, level, level, 8)[0]
    wl, hl = W >> level, H >> level
    nop = 2 if mode == 1 else 1
    rng = np.random.default_rng(0)
    fl = (rng.standard_normal((hl, wl, nop)) * 0.7).astype(np.float32)
    for camlr in ((0,) if mode == 1 else (0, 1)):
        f = fl if mode == 1 else (-np.abs(fl) if camlr == 0 else np.abs(fl))
        want = port.varref_level(q, a, b, wl, hl, level, camlr, f)
        got = gpu.varref_level(p, a, b, wl, hl, level, camlr, f)
        assert np.abs(want - f).max() > 0.1          # the refinement did something
        assert np.array_equal(got, want), "max diff %g" % np.abs(got - want).max()


@pytest.mark.parametrize("solverit,innerit,omega", [(1, 1, 1.0), (5, 2, 1.9), (3, 1, 1.6)])
def test_varref_sweep_counts(gpu, port, solverit, innerit, omega):
    A, B, _ = make_pair(120, 200, seed=6)
    q = base_orc(1, 1, sc_f=1, sc_l=1, tv_solverit=solverit, tv_innerit=innerit, tv_sor=omega)
    a, b = port.pyramid_level(A, 1, 1, 8)[0], port.pyramid_level(B, 1, 1, 8)[0]
    fl = np.full((60, 100, 2), 0.3, np.float32)
    assert np.array_equal(gpu.varref_level(to_dis(gpu, q), a, b, 100, 60, 1, 0, fl), port.varref_level(q, a, b, 100, 60, 1, 0, fl))


def test_varref_tall_level_uses_band_mode(gpu, port):
    """h * solverit > 1024 threads: sweeps run band by band, still the lexicographic order."""
    A, B, _ = make_pair(400, 96, seed=8)
    q = base_orc(1, 1, sc_f=0, sc_l=0, tv_solverit=3)
    a, b = port.pyramid_level(A, 0, 0, 8)[0], port.pyramid_level(B, 0, 0, 8)[0]
    fl = np.zeros((400, 96, 2), np.float32)
    fl[..., 0] = 1.5
    assert np.array_equal(gpu.varref_level(to_dis(gpu, q), a, b, 96, 400, 0, 0, fl), port.varref_level(q, a, b, 96, 400, 0, 0, fl))


@pytest.mark.parametrize("mode,ch", [(1, 1), (1, 3), (2, 1), (2, 3)])
@pytest.mark.parametrize("w0,h0,level,solverit", [(1024, 436, 3, 3), (500, 218, 2, 3), (260, 128, 1, 5), (250, 120, 2, 2), (1920, 1080, 4, 3), (300, 180, 1, 1),
                                                  (500, 436, 1, 3), (640, 400, 0, 2), (333, 300, 0, 1), (256, 120, 2, 3), (1024, 436, 4, 3),
                                                  (500, 218, 2, 2), (40, 300, 2, 3), (1280, 720, 3, 1)])
def test_tv_paths_agree_bit_for_bit(gpu, port, monkeypatch, mode, ch, w0, h0, level, solverit):
    """The schedules of the lexicographic SOR — one CTA per pair (fused), one warp per pair with shuffles and cp.async
    row streams (split; levels taller than 96 rows: one warp per band of 32 rows, bands chained through tagged boundary
    rows), general wavefront (old), system assembled inside the SOR kernel, and the team schedule (one warp per band in
    one CTA per pair, the automatic choice for 33 - 96 rows: two- and three-band levels with 1, 2 and 3 sweeps, a level
    ten pixels wide) — are the same iteration."""
    A, B, _ = make_pair(h0, w0, channels=ch, mode="OF" if mode == 1 else "DE", seed=5)
    q = base_orc(mode, ch, sc_f=level, sc_l=level, tv_solverit=solverit)
    W, H, _, _ = port.pad_dims(w0, h0, level)
    a, b = port.pyramid_level(A, level, level, 8)[0], port.pyramid_level(B, level, level, 8)[0]
    wl, hl = W >> level, H >> level
    nop = 2 if mode == 1 else 1
    fl = (np.random.default_rng(1).standard_normal((hl, wl, nop)) * 0.5).astype(np.float32)
    if mode == 2:
        fl = -np.abs(fl)
    want = port.varref_level(q, a, b, wl, hl, level, 0, fl)
    try:
        for sched in (gpu.TV_FUSED, gpu.TV_SPLIT, gpu.TV_GENERAL, gpu.TV_SOLVEF, gpu.TV_TEAM):
            gpu.set_tuning(tv_schedule=sched)
            got = gpu.varref_level(to_dis(gpu, q), a, b, wl, hl, level, 0, fl)
            assert np.array_equal(got, want), (sched, float(np.abs(got - want).max()))
    finally:
        gpu.set_tuning()


@pytest.mark.parametrize("mode,ch", [(1, 1), (2, 1), (1, 3)])
def test_varref_redblack_matches_its_oracle(gpu, port, mode, ch):
    """DISB200_SOR_REDBLACK is not the reference's order; the oracle restates it (orc_set_sor_order)."""
    w0, h0, level = 500, 218, 1
    A, B, _ = make_pair(h0, w0, channels=ch, mode="OF" if mode == 1 else "DE", seed=3)
    q = base_orc(mode, ch, sc_f=level, sc_l=level)
    W, H, _, _ = port.pad_dims(w0, h0, level)
    a, b = port.pyramid_level(A, level, level, 8)[0], port.pyramid_level(B, level, level, 8)[0]
    wl, hl = W >> level, H >> level
    fl = np.full((hl, wl, 2 if mode == 1 else 1), -0.4, np.float32)
    lex = port.varref_level(q, a, b, wl, hl, level, 0, fl)
    try:
        port.set_sor_order(1)
        want = port.varref_level(q, a, b, wl, hl, level, 0, fl)
    finally:
        port.set_sor_order(0)
    got = gpu.varref_level(to_dis(gpu, q, sor_mode=1), a, b, wl, hl, level, 0, fl)
    assert np.array_equal(got, want), float(np.abs(got - want).max())
    assert not np.array_equal(want, lex)                      # a different iterate ...
    assert np.abs(want - lex).mean() < 0.1                    # ... of the same problem


# ------------------------------------------------------------------------------ OFClass level (drop-in boundary)
@pytest.mark.parametrize("name", sorted(REF_CASES))
def test_ofclass_against_reference_golden(gpu, port, ref_golden, name):
    """Same padded host pyramids in, low-resolution flow out, against outputs of the reference's own build."""
    mode, noc, h0, w0, kw = REF_CASES[name]
    A, B = ref_golden[name + "_A"], ref_golden[name + "_B"]
    q = base_orc(mode, noc, **kw)
    W, H, _, _ = port.pad_dims(w0, h0, q.sc_f)
    PA, PB = port.pyramid(A, q.sc_f, q.sc_l, q.psz), port.pyramid(B, q.sc_f, q.sc_l, q.psz)
    got = gpu.ofclass(to_dis(gpu, q), PA, PB, W, H)
    m, mx = epe(got, ref_golden[name + "_low"])
    assert m <= MEAN_TOL and mx <= (MAX_TOL_DENSE if q.psz == 12 else 0.05), (m, mx)


def test_ofclass_initflow(gpu, port):
    A, B, _ = make_pair(96, 152, seed=11)      # W, H divisible by 2^(sc_f + 1): what initflow requires (run_dense.cpp:301)
    q = base_orc(1, 1, sc_f=2, sc_l=1)
    W, H, _, _ = port.pad_dims(152, 96, 2)
    PA, PB = port.pyramid(A, 2, 1, 8), port.pyramid(B, 2, 1, 8)
    init = np.zeros((H >> 3, W >> 3, 2), np.float32)
    init[..., 0] = 0.4
    init[..., 1] = -0.2
    want = port.ofclass(q, PA, PB, W, H, initflow=init)
    got = gpu.ofclass(to_dis(gpu, q), PA, PB, W, H, initflow=init)
    assert not np.array_equal(want, port.ofclass(q, PA, PB, W, H))
    # the handle-based variant (no allocation per call) gives the same bits, with and without initflow, call after call
    with gpu.Handle(to_dis(gpu, q), W, H, max_batch=1) as hd:
        cold = hd.ofclass(PA, PB)
        assert np.array_equal(hd.ofclass(PA, PB, initflow=init), got)
        assert np.array_equal(hd.ofclass(PA, PB), cold) and np.array_equal(cold, gpu.ofclass(to_dis(gpu, q), PA, PB, W, H))
    m, mx = epe(got, want)
    assert m <= MEAN_TOL and mx <= 0.05


# ------------------------------------------------------------------------------ whole driver path, host buffers
CONFIGS = [
    # id, mode, ch, op, w0, h0, overrides, max tolerance
    ("C1_of_int_op2", 1, 1, 2, 1024, 436, {}, MAX_TOL_FINE),
    ("of_int_op1", 1, 1, 1, 1024, 436, {}, MAX_TOL_FINE),
    ("of_int_op3", 1, 1, 3, 1024, 436, {}, MAX_TOL_DENSE),
    ("of_rgb_op2", 1, 3, 2, 1024, 436, {}, MAX_TOL_FINE),
    ("of_rgb_op3", 1, 3, 3, 640, 272, {}, MAX_TOL_DENSE),
    ("C3_de_int_op2", 2, 1, 2, 1280, 720, {}, MAX_TOL_FINE),
    ("de_rgb_op2", 2, 3, 2, 1280, 720, {}, MAX_TOL_FINE),
    ("of_int_l1", 1, 1, 2, 1024, 436, dict(costfct=1), MAX_TOL_FINE),
    ("of_int_huber", 1, 1, 2, 1024, 436, dict(costfct=2), MAX_TOL_FINE),
    ("of_int_fb", 1, 1, 2, 1024, 436, dict(usefbcon=1), MAX_TOL_FINE),
    # NCC has no reference semantics: the oracle restates the definition of DESIGN.md "NCC"
    ("of_int_ncc", 1, 1, 2, 1024, 436, dict(costfct=10), 0.05),
    ("of_int_ncc_fb", 1, 1, 2, 640, 272, dict(costfct=10, usefbcon=1), 0.05),
    ("of_rgb_l1_fb", 1, 3, 2, 640, 272, dict(costfct=1, usefbcon=1), 0.05),
    ("de_int_ncc", 2, 1, 2, 1280, 720, dict(costfct=10), 0.05),
    ("de_int_fb_op3", 2, 1, 3, 640, 360, dict(usefbcon=1), MAX_TOL_DENSE),
    ("of_int_early_stop", 1, 1, 2, 1024, 436, dict(min_iter=2, max_iter=30, res_thresh=0.4), 0.05),
    ("of_int_nonorm", 1, 1, 2, 1024, 436, dict(patnorm=0), MAX_TOL_FINE),
    ("of_int_psz16", 1, 1, 2, 640, 272, dict(patchsz=16, patove=0.5, sc_f=3, sc_l=1), 0.05),
    ("of_int_psz4", 1, 1, 2, 320, 136, dict(patchsz=4, patove=0.5, sc_f=3, sc_l=2), 0.05),
    # every patch size the reference accepts (oflow.cpp:90-91): even sizes without a specialised kernel, and ODD sizes,
    # where the reference samples (psz - 1)^2 positions into vectors of psz^2 entries and LossComputeErrorImage skips
    # the last novals % 4 of them (patch.cpp:226-232,316-317) - oracle pinned against oracle/_ref on the same cases
    ("of_int_psz6", 1, 1, 2, 640, 272, dict(patchsz=6, patove=0.5, sc_f=3, sc_l=1), 0.05),
    ("of_int_psz9", 1, 1, 2, 640, 272, dict(patchsz=9, patove=0.5, sc_f=3, sc_l=1), 0.05),
    ("of_rgb_psz5", 1, 3, 2, 320, 136, dict(patchsz=5, patove=0.4, sc_f=3, sc_l=1), 0.05),
    ("de_int_psz7_huber", 2, 1, 2, 640, 360, dict(patchsz=7, patove=0.4, sc_f=3, sc_l=1, costfct=2), 0.05),
    ("of_int_psz10_fb", 1, 1, 2, 640, 272, dict(patchsz=10, patove=0.6, sc_f=3, sc_l=1, usefbcon=1), 0.05),
    ("of_rgb_psz15_fb", 1, 3, 2, 320, 136, dict(patchsz=15, patove=0.7, sc_f=2, sc_l=0, usefbcon=1), 0.05),
    ("of_int_psz2", 1, 1, 2, 160, 96, dict(patchsz=2, patove=0.0, sc_f=2, sc_l=0), 0.05),
]


@pytest.mark.parametrize("cfg", CONFIGS, ids=[c[0] for c in CONFIGS])
def test_run_dense_parity(gpu, port, cfg):
    _, mode, ch, op, w0, h0, kw, max_tol = cfg
    A, B, gt = make_pair(h0, w0, channels=ch, mode="OF" if mode == 1 else "DE", seed=0)
    p = gpu.operating_point(op, w0, mode, ch).copy(verbosity=0, **kw)
    want, want_low = port.run_dense(to_orc(p), A, B, want_lowres=True)
    with gpu.Handle(p, w0, h0, max_batch=3) as hd:
        out = hd.run_host(np.stack([A, B, A]), np.stack([B, A, B]))
        low = hd.fetch_flow(0, p.sc_l, pair=2)
        assert hd.launch_count() > 0
    m, mx = epe(out[0], want)
    assert m <= MEAN_TOL and mx <= max_tol, "EPE(GPU,CPU) mean %g max %g" % (m, mx)
    assert np.array_equal(out[0], out[2])                     # batch position does not matter
    assert epe(low, want_low)[0] <= MEAN_TOL
    # the flow is a flow: error against the synthetic ground truth is what the CPU reference gets
    assert abs(epe(out[0], gt)[0] - epe(want, gt)[0]) < 5e-3
    # swapped pair (B -> A) is a different problem and must not leak from its neighbours
    want_swapped = port.run_dense(to_orc(p), B, A)
    assert epe(out[1], want_swapped)[0] <= MEAN_TOL


@pytest.mark.parametrize("case", TINY_CASES)
def test_tiny_levels_and_zero_sweeps(gpu, port, case):
    """Levels down to one pixel wide (sor_coupled's point-wise fallback, solver.c:86-89), 1 x 1 levels without
    refinement, tv_solverit = 0: same cases on which the oracle is pinned against oracle/_ref in the CPU suite."""
    mode, noc, w0, h0, sc_f, sc_l, kw = case
    A, B, _ = make_pair(h0, w0, channels=noc, mode="OF" if mode == 1 else "DE", seed=1)
    q = base_orc(mode, noc, sc_f=sc_f, sc_l=sc_l, **kw)
    want = port.run_dense(q, A, B)
    with gpu.Handle(to_dis(gpu, q), w0, h0, max_batch=2) as hd:
        out = hd.run_host(np.stack([A, A]), np.stack([B, B]))
    assert np.isfinite(out).all()
    m, mx = epe(out[0], want)
    assert m <= MEAN_TOL and mx <= 0.05, "EPE(GPU,CPU) mean %g max %g" % (m, mx)
    assert np.array_equal(out[0], out[1])


def test_one_row_level_with_refinement_is_refused(gpu):
    """The reference segfaults on a level one row high when the refinement is on (probe in DESIGN.md): refused, not guessed."""
    q = base_orc(1, 1, sc_f=4, sc_l=0)
    with pytest.raises(gpu.DisError):
        gpu.Handle(to_dis(gpu, q), 64, 16, max_batch=1)


@pytest.mark.parametrize("mode,noc,psz,ov", [(1, 1, 8, .4), (1, 3, 8, .4), (2, 1, 12, .75), (1, 1, 5, .4), (2, 3, 9, .5)])
@pytest.mark.parametrize("spread", ["regular", "collapsed"])
def test_aggregate_dense_cut(gpu, port, mode, noc, psz, ov, spread):
    """AggregateFlowDense (patchgrid.cpp:213-397) on hand-made patch results, with forward-backward merging.
    "collapsed": every patch of the complementary grid ends inside one patch-sized cell (a contracting / occluded
    region) - hundreds of entries per cell, which the CSR binning must take like any other distribution."""
    w_l, h_l = 150, 90
    q = base_orc(mode, noc, psz=psz, patove=ov)
    p = to_dis(gpu, q)
    n = gpu.patch_count(p, w_l, h_l)
    nop, nv = p.nop, noc * psz * psz
    rng = np.random.default_rng(7)
    p_iter = rng.standard_normal((n, nop)).astype(np.float32) * 3
    pweight = (rng.random((n, nv)) * 9).astype(np.float32)
    cg_p = rng.standard_normal((n, nop)).astype(np.float32) * 3
    cg_w = (rng.random((n, nv)) * 9).astype(np.float32)
    if spread == "regular":
        cg_pt = np.stack([rng.random(n) * (w_l + 20) - 10, rng.random(n) * (h_l + 20) - 10], -1).astype(np.float32)
    else:
        cg_pt = np.stack([60 + rng.random(n) * 5, 40 + rng.random(n) * 5], -1).astype(np.float32)
    for cg in (None, (cg_p, cg_pt, cg_w)):
        want = port.aggregate_dense(q, w_l, h_l, p_iter, pweight, cg)
        got = gpu.aggregate_dense(p, w_l, h_l, p_iter, pweight, cg)
        if cg is None:
            assert np.array_equal(got, want)      # the gather adds the covering patches in the reference's order
        else:
            assert np.allclose(got, want, rtol=2e-4, atol=2e-4), float(np.abs(got - want).max())


@pytest.mark.parametrize("mode,ch,op,w0,h0", [(1, 1, 2, 1024, 436), (1, 3, 2, 640, 272), (2, 1, 2, 1280, 720), (1, 1, 3, 512, 218)])
def test_against_the_compiled_reference_directly(gpu, port, mode, ch, op, w0, h0):
    """No indirection: the CUDA path against oracle/_ref (the reference's own sources, compiled unmodified by
    oracle/Makefile; the .so travels to the GPU box) on the same pyramids - OFClass cut and whole driver path."""
    if not ob.have_ref(mode, ch):
        pytest.skip("oracle/_ref did not travel with this snapshot")
    ref = ob.Ref(mode, ch)
    A, B, _ = make_pair(h0, w0, channels=ch, mode="OF" if mode == 1 else "DE", seed=4)
    p = gpu.operating_point(op, w0, mode, ch).copy(verbosity=0)
    q = to_orc(p)
    W, H, padl, padt = port.pad_dims(w0, h0, p.sc_f)
    PA = port.pyramid(A, p.sc_f, p.sc_l, p.patchsz)
    PB = port.pyramid(B, p.sc_f, p.sc_l, p.patchsz)
    want_low = ref.ofclass(q, PA, PB, W, H)
    got_low = gpu.ofclass(p, PA, PB, W, H)
    with gpu.Handle(p, w0, h0, max_batch=1) as hd:
        got = hd.run_host(A[None], B[None])[0]
    want = port.upsample_crop(want_low, p.sc_l, padl, padt, w0, h0)
    tol = MAX_TOL_FINE if op == 2 else MAX_TOL_DENSE
    for g, w in ((got_low, want_low), (got, want)):
        m, mx = epe(g, w)
        assert m <= MEAN_TOL and mx <= tol, "EPE(GPU, compiled reference) mean %g max %g" % (m, mx)
    if op == 2 and ch == 1:
        assert np.array_equal(got_low, want_low) and np.array_equal(got, want)   # grey op2: bit-identical to the reference


def test_c2_rgb_op4_full_size(gpu, port):
    """BASELINE.json configs[1]: run_OF_RGB operating point 4 (128 iterations, psz 12, overlap .75, scales 5..0)."""
    A, B, gt = make_pair(436, 1024, channels=3, seed=0)
    p = gpu.operating_point(4, 1024, 1, 3).copy(verbosity=0)
    assert (p.sc_f, p.sc_l, p.max_iter, p.patchsz) == (5, 0, 128, 12)
    want = port.run_dense(to_orc(p), A, B)
    with gpu.Handle(p, 1024, 436, 1) as hd:
        out = hd.run_host(A[None], B[None])[0]
    m, mx = epe(out, want)
    assert m <= MEAN_TOL and mx <= MAX_TOL_DENSE, (m, mx)
    assert abs(epe(out, gt)[0] - epe(want, gt)[0]) < 5e-3


def test_c4_4k_cost_variants_with_fb_consistency(gpu, port):
    """BASELINE.json configs[3] on ONE GPU: 3840x2160, forward-backward consistency, L1 / Huber / NCC, explicit
    parameters `7 2 16 16 .05 .95 0 8 .40 1 1 {1,2,10} 1 10 10 5 1 3 1.6 0` (lighter line)."""
    A, B, gt = make_pair(2160, 3840, seed=0)
    base = gpu.operating_point(2, 3840).copy(verbosity=0, sc_f=7, sc_l=2, max_iter=16, min_iter=16, usefbcon=1)
    for cost in (1, 2, 10):
        p = base.copy(costfct=cost)
        want = port.run_dense(to_orc(p), A, B)
        with gpu.Handle(p, 3840, 2160, 1) as hd:
            out = hd.run_host(A[None], B[None])[0]
        m, mx = epe(out, want)
        assert m <= MEAN_TOL and mx <= 0.25, (cost, m, mx)
        assert epe(out, gt)[0] < 0.5


def test_c4_heavy_line_reduced_size(gpu, port):
    """The heavy C4 line of SURVEY.md 8(d) (`7 1 16 16 .05 .95 0 12 .75 1 1 1 ...`: 12 x 12 patches at overlap .75 down to
    level 1, forward-backward consistency, L1) against the oracle, on a 1920x1080 frame so that the CPU side stays within
    seconds; the full 3840x2160 pair is in bench.py --workload C4h (parity object: mean 7.5e-6 px) and in
    tests/test_tiled_multi_gpu.py."""
    A, B, gt = make_pair(1080, 1920, seed=0)
    p = gpu.operating_point(2, 1920).copy(verbosity=0, sc_f=6, sc_l=1, max_iter=16, min_iter=16, patchsz=12, patove=0.75, usefbcon=1, costfct=1)
    want = port.run_dense(to_orc(p), A, B)
    with gpu.Handle(p, 1920, 1080, 1) as hd:
        out = hd.run_host(A[None], B[None])[0]
    m, mx = epe(out, want)
    assert m <= MEAN_TOL and mx <= MAX_TOL_DENSE, (m, mx)
    assert abs(epe(out, gt)[0] - epe(want, gt)[0]) < 5e-3


def test_c5_full_size_video_against_the_oracle(gpu, port):
    """BASELINE.json configs[4] at its real frame size: consecutive 1920x1080 frames through the sequence entry point
    (pyramids shared between pairs), every pair against the CPU oracle - op2 grey is bit-identical."""
    from of_dis_b200.synth import make_sequence
    fr = np.stack(make_sequence(1080, 1920, 4, seed=0))
    p = gpu.operating_point(2, 1920).copy(verbosity=0)
    with gpu.Handle(p, 1920, 1080, max_batch=3) as hd:
        got = hd.run_sequence(fr)
    for i in range(3):
        want = port.run_dense(to_orc(p), fr[i], fr[i + 1])
        assert np.array_equal(got[i], want), (i, epe(got[i], want))


def test_run_list_and_chunking(gpu, port):
    """disb200_run_batch (pointer lists) and n > max_batch chunking give the same answers."""
    pairs = [make_pair(100, 180, seed=s)[:2] for s in range(5)]
    p = gpu.operating_point(2, 180).copy(verbosity=0, sc_f=2, sc_l=1)
    As, Bs = [a for a, _ in pairs], [b for _, b in pairs]
    with gpu.Handle(p, 180, 100, max_batch=2) as hd:
        outs = hd.run_list(As, Bs)
        outs2 = hd.run_host(np.stack(As), np.stack(Bs))
    for i in range(5):
        assert np.array_equal(outs[i], outs2[i])
        assert epe(outs[i], port.run_dense(to_orc(p), As[i], Bs[i]))[0] <= MEAN_TOL


def test_redblack_whole_path(gpu, port):
    """Throughput mode end to end: its own EPE against the CPU reference and against the ground truth."""
    A, B, gt = make_pair(436, 1024, seed=0)
    p = gpu.operating_point(2, 1024).copy(verbosity=0)
    want = port.run_dense(to_orc(p), A, B)
    with gpu.Handle(p.copy(sor_mode=1), 1024, 436, 1) as hd:
        out = hd.run_host(A[None], B[None])[0]
    m, mx = epe(out, want)
    print("red-black vs lexicographic CPU: mean %.4f max %.3f px; vs GT %.4f (CPU %.4f)" % (m, mx, epe(out, gt)[0], epe(want, gt)[0]))
    assert 1e-4 < m < 0.15                                    # not the parity iterate, but the same flow
    assert abs(epe(out, gt)[0] - epe(want, gt)[0]) < 0.05


def test_cli_matches_library(gpu, tmp_path):
    """run_OF_INT / run_DE_INT (reference argv grammar, PGM in, .flo / .pfm out) against the batch entry point."""
    import subprocess
    exe_dir = os.path.join(ROOT, "of_dis_b200", "bin")
    for mode, exe, ext in ((1, "run_OF_INT", "flo"), (2, "run_DE_INT", "pfm"), (1, "run_OF_GRAD", "flo")):
        A, B, _ = make_pair(120, 200, mode="OF" if mode == 1 else "DE", seed=2)
        for name, img in (("a.pgm", A), ("b.pgm", B)):
            with open(tmp_path / name, "wb") as f:
                f.write(b"P5\n200 120\n255\n" + img.tobytes())
        out = tmp_path / ("out." + ext)
        r = subprocess.run([os.path.join(exe_dir, exe), str(tmp_path / "a.pgm"), str(tmp_path / "b.pgm"), str(out), "2"],
                           capture_output=True, text=True)
        assert r.returncode == 0, r.stderr
        assert "O.Flow Run-Time" in r.stdout
        p = gpu.operating_point(2, 200, mode, 1).copy(gradmag=int(exe.endswith("GRAD")))
        # operating points print at verbosity 2 (run_dense.cpp:234): one "TIME (Sc: l, #p: n, ...)" line per scale (oflow.cpp:297)
        sc_lines = [l for l in r.stdout.splitlines() if l.startswith("TIME (Sc: ")]
        assert [int(l.split(",")[0].split(":")[1]) for l in sc_lines] == list(range(p.sc_f, p.sc_l - 1, -1)), r.stdout
        with gpu.Handle(p, 200, 120, 1) as hd:
            want = hd.run_host(A[None], B[None])[0]
        raw = open(out, "rb").read()
        if mode == 1:
            assert raw[:4] == b"PIEH" and np.frombuffer(raw[4:12], np.int32).tolist() == [200, 120]
            got = np.frombuffer(raw[12:], np.float32).reshape(120, 200, 2)
        else:
            hdr = b"Pf\n200 120\n-1.000000\n"
            assert raw.startswith(hdr)
            got = -np.frombuffer(raw[len(hdr):], np.float32).reshape(120, 200)[::-1, :, None]
        assert np.array_equal(got, want)
    # malformed argument lists are errors, not out-of-bounds reads (run_dense.cpp:272-292)
    assert subprocess.run([os.path.join(exe_dir, "run_OF_INT"), "x", "y", "z", "1", "2", "3"], capture_output=True).returncode != 0


def _write_png_grey(path, img):
    """Minimal 8-bit grey PNG (filter 0), enough to feed the CLIs."""
    import struct
    import zlib

    def chunk(tag, body):
        return struct.pack(">I", len(body)) + tag + body + struct.pack(">I", zlib.crc32(tag + body) & 0xFFFFFFFF)
    h, w = img.shape
    raw = b"".join(b"\0" + img[y].tobytes() for y in range(h))
    with open(path, "wb") as f:
        f.write(b"\x89PNG\r\n\x1a\n" + chunk(b"IHDR", struct.pack(">IIBBBBB", w, h, 8, 0, 0, 0, 0))
                + chunk(b"IDAT", zlib.compress(raw)) + chunk(b"IEND", b""))


def _read_flo(path, w, h):
    raw = open(path, "rb").read()
    assert raw[:4] == b"PIEH" and np.frombuffer(raw[4:12], np.int32).tolist() == [w, h]
    return np.frombuffer(raw[12:], np.float32).reshape(h, w, 2)


def test_batch_cli_png_in_flo_out_matches_library(gpu, tmp_path):
    """disb200_batch: PNG decode -> batched flow (batches of 2, the last one ragged) -> writer pool, and the
    --video form (one pyramid per frame); both bit-identical with the library calls they wrap."""
    import subprocess
    from of_dis_b200.synth import make_sequence
    exe = os.path.join(ROOT, "of_dis_b200", "bin", "disb200_batch")
    w0, h0 = 200, 120
    pairs = [make_pair(h0, w0, seed=s)[:2] for s in range(5)]
    lines = []
    for i, (A, B) in enumerate(pairs):
        _write_png_grey(tmp_path / ("a%d.png" % i), A)
        _write_png_grey(tmp_path / ("b%d.png" % i), B)
        lines.append("%s %s %s" % (tmp_path / ("a%d.png" % i), tmp_path / ("b%d.png" % i), tmp_path / ("o%d.flo" % i)))
    (tmp_path / "list.txt").write_text("# five pairs\n" + "\n".join(lines) + "\n")
    r = subprocess.run([exe, "--batch", "2", "--threads", "3", str(tmp_path / "list.txt"), "2"], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    p = gpu.operating_point(2, w0, 1, 1).copy(verbosity=0)
    with gpu.Handle(p, w0, h0, 5) as hd:
        want = hd.run_host(np.stack([a for a, _ in pairs]), np.stack([b for _, b in pairs]))
    for i in range(5):
        assert np.array_equal(_read_flo(tmp_path / ("o%d.flo" % i), w0, h0), want[i]), i
    # video: 6 frames -> 5 flows, chunks of 3 pairs
    fr = make_sequence(h0, w0, 6, seed=1)
    vl = []
    for i, f in enumerate(fr):
        _write_png_grey(tmp_path / ("f%d.png" % i), f)
        vl.append("%s %s" % (tmp_path / ("f%d.png" % i), tmp_path / ("v%d.flo" % i)))
    (tmp_path / "video.txt").write_text("\n".join(vl) + "\n")
    r = subprocess.run([exe, "--video", "--batch", "3", str(tmp_path / "video.txt")], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    with gpu.Handle(p, w0, h0, 5) as hd:
        want = hd.run_sequence(np.stack(fr))
    for i in range(5):
        assert np.array_equal(_read_flo(tmp_path / ("v%d.flo" % i), w0, h0), want[i]), i
    assert not (tmp_path / "v5.flo").exists()
    # a missing input is an error, not a silent skip
    (tmp_path / "bad.txt").write_text("%s %s %s\n" % (tmp_path / "a0.png", tmp_path / "nope.png", tmp_path / "x.flo"))
    assert subprocess.run([exe, str(tmp_path / "bad.txt")], capture_output=True).returncode != 0


def test_band_sor_many_pairs_matches_general_wavefront(gpu, monkeypatch):
    """Several pairs x several bands in one sor_band_kernel launch (ticket-ordered CTAs, one exchange region per
    pair): the whole path with tall levels (400 and 200 rows) equals the run that forces the general wavefront,
    for every pair, and repeated runs reuse the exchange rows under a new epoch without cross-talk."""
    w0, h0, n = 208, 400, 5
    pairs = [make_pair(h0, w0, seed=20 + s)[:2] for s in range(n)]
    A, B = np.stack([a for a, _ in pairs]), np.stack([b for _, b in pairs])
    p = gpu.operating_point(2, w0).copy(verbosity=0, sc_f=2, sc_l=0, usefbcon=1)
    try:
        gpu.set_tuning(tv_schedule=gpu.TV_GENERAL)
        with gpu.Handle(p, w0, h0, n) as hd:
            want = hd.run_host(A, B)
    finally:
        gpu.set_tuning()
    with gpu.Handle(p, w0, h0, n) as hd:
        got1 = hd.run_host(A, B)
        got2 = hd.run_host(A[::-1].copy(), B[::-1].copy())[::-1]
        got3 = hd.run_host(A[:2], B[:2])
    assert np.array_equal(got1, want)
    assert np.array_equal(got2, want)
    assert np.array_equal(got3, want[:2])


def test_ncc_is_gain_invariant(gpu):
    """Second frame darkened by 30 % (+ offset): the NCC flow barely moves, the L2 flow does."""
    A, B, gt = make_pair(218, 512, seed=4)
    B2 = np.clip(B.astype(np.float32) * 0.7 + 20, 0, 255).astype(np.uint8)
    res = {}
    for cost in (0, 10):
        p = gpu.operating_point(2, 512).copy(verbosity=0, costfct=cost, usetvref=0)
        with gpu.Handle(p, 512, 218, 2) as hd:
            out = hd.run_host(np.stack([A, A]), np.stack([B, B2]))
        res[cost] = (epe(out[0], gt)[0], epe(out[1], gt)[0])
    assert res[10][1] < res[10][0] + 0.1          # NCC: same quality under the gain change
    assert res[10][1] < res[0][1]                 # and better than L2 on the darkened frame


def test_video_sequence_reuses_pyramids_and_matches_explicit_pairs(gpu):
    """C5-style front end: n + 1 frames -> n flows, bit-identical to the explicit (i, i+1) pairs, chunked too."""
    from of_dis_b200.synth import make_pair
    frames = np.stack([make_pair(136, 240, seed=s)[0] for s in range(7)])
    for kw in (dict(), dict(usefbcon=1)):
        p = gpu.operating_point(2, 240).copy(verbosity=0, sc_f=3, sc_l=1, **kw)
        with gpu.Handle(p, 240, 136, max_batch=4) as hd:
            want = hd.run_host(frames[:-1].copy(), frames[1:].copy())
            l_pairs = hd.launch_count()
            got = hd.run_sequence(frames)              # 6 pairs, max_batch 4 -> two chunks
            got3 = hd.run_sequence(frames[:4].copy())  # one chunk
            assert hd.launch_count() > 0 and l_pairs > 0
        assert np.array_equal(got, want)
        assert np.array_equal(got3, want[:3])


@pytest.mark.parametrize("mode,ch", [(1, 1), (2, 1), (1, 3)])
def test_video_sequence_temporal_warm_start(gpu, port, mode, ch):
    """disb200_run_sequence_warm_*: pair i starts from pair i-1's result through OFClass's `initflow` (oflow.cpp:213-217).
    Checked against the CPU oracle called pair by pair with the same initflow (box mean of its own previous level-sc_l
    flow, the INTER_AREA rule of the reference's disabled caller, run_dense.cpp:370-373)."""
    from of_dis_b200.synth import make_sequence
    h0, w0 = 128, 240          # padded dimensions divisible by 2^(sc_f + 1), as initflow requires (run_dense.cpp:301)
    fr = make_sequence(h0, w0, 5, seed=1)
    frames = np.stack(fr) if ch == 1 else np.stack([np.stack([f, np.roll(f, 3, 0), np.roll(f, 5, 1)], -1) for f in fr])
    q = base_orc(mode, ch, sc_f=3, sc_l=1)
    p = to_dis(gpu, q)
    W, H, padl, padt = port.pad_dims(w0, h0, q.sc_f)
    k = 1 << (q.sc_f + 1 - q.sc_l)
    hi, wi = (H >> q.sc_f) // 2, (W >> q.sc_f) // 2
    nop = p.nop
    want, prev = [], None
    for i in range(len(frames) - 1):
        PA = port.pyramid(frames[i], q.sc_f, q.sc_l, q.psz)
        PB = port.pyramid(frames[i + 1], q.sc_f, q.sc_l, q.psz)
        init = None
        if prev is not None:
            init = (prev[:hi * k, :wi * k].reshape(hi, k, wi, k, nop).astype(np.float64).sum(axis=(1, 3)) / (k * k) / k).astype(np.float32)
        prev = port.ofclass(q, PA, PB, W, H, initflow=init)
        want.append(port.upsample_crop(prev, q.sc_l, padl, padt, w0, h0))
    want = np.stack(want)
    with gpu.Handle(p, w0, h0, max_batch=3) as hd:        # 4 pairs, capacity 3: two chunks, the warm state carries over
        got = hd.run_sequence_warm(frames)
        cold = hd.run_sequence(frames)
        assert hd.launch_count() > 0
    assert np.array_equal(got[0], cold[0])                # pair 0 starts cold
    if mode == 1:
        assert not np.array_equal(got[1:], cold[1:])      # the others start somewhere else
    m, mx = epe(got, want)
    assert m <= MEAN_TOL and mx <= 0.05, "EPE(GPU,CPU) mean %g max %g" % (m, mx)
    # odd coarsest level: initflow would be read out of bounds (by the reference too) - refused
    with gpu.Handle(p, w0, 136, max_batch=3) as hd:
        with pytest.raises(gpu.DisError):
            hd.run_sequence_warm(np.zeros((3, 136, w0) + ((ch,) if ch > 1 else ()), np.uint8))


@pytest.mark.parametrize("mode,ch,op,w0,h0", [(1, 1, 2, 1024, 436), (2, 1, 2, 640, 360), (1, 3, 2, 320, 136), (1, 1, 4, 150, 90)])
def test_lowres_entry_points_return_the_ofclass_outflow(gpu, port, mode, ch, op, w0, h0):
    """disb200_run_batch_host_lowres / run_sequence_host_lowres: the level-sc_l flow (what OFClass writes to outflow,
    oflow.h:105), chunked like the full-resolution calls; upsampling it with the oracle's restatement of
    run_dense.cpp:407-415 gives exactly the full-resolution result.  op4 on 150x90 has sc_l = 0 (padded level larger
    than the cropped output)."""
    A, B, _ = make_pair(h0, w0, channels=ch, mode="OF" if mode == 1 else "DE", seed=6)
    p = gpu.operating_point(op, w0, mode, ch).copy(verbosity=0)
    W, H, padl, padt = port.pad_dims(w0, h0, p.sc_f)
    As, Bs = np.stack([A, B, A, B, A]), np.stack([B, A, B, A, B])
    with gpu.Handle(p, w0, h0, max_batch=2) as hd:          # 5 pairs, max_batch 2 -> three chunks
        assert hd.lowres_dims() == (H >> p.sc_l, W >> p.sc_l, p.nop)
        full = hd.run_host(As, Bs)
        low = hd.run_host_lowres(As, Bs)
        assert hd.launch_count() > 0
        assert np.array_equal(low[0], hd.fetch_flow(0, p.sc_l, pair=0))    # last chunk holds pair 4 == pair 0
        assert np.array_equal(hd.run_host(As, Bs), full)                   # the mode does not stick to the handle
        frames = np.stack([A, B, A, B])
        seq_low = hd.run_sequence_lowres(frames)
    assert low.shape == (5, H >> p.sc_l, W >> p.sc_l, p.nop)
    for i in range(5):
        assert np.array_equal(port.upsample_crop(low[i], p.sc_l, padl, padt, w0, h0), full[i]), i
    assert np.array_equal(seq_low, low[:3])
    want_low = port.run_dense(to_orc(p), A, B, want_lowres=True)[1]
    assert epe(low[0], want_low)[0] <= MEAN_TOL


def test_identical_frames_give_exactly_zero(gpu):
    A, _, _ = make_pair(218, 512, seed=1)
    p = gpu.operating_point(2, 512).copy(verbosity=0)
    with gpu.Handle(p, 512, 218, 1) as hd:
        assert np.all(hd.run_host(A[None], A[None]) == 0)


def test_constant_frames(gpu, port):
    """Zero-texture patches: mares == 0 -> immediate convergence; Hessian regulariser (patch.cpp:76-84)."""
    A = np.full((96, 160), 77, np.uint8)
    B = A.copy()
    B[:, 80:] = 90
    p = gpu.operating_point(2, 160).copy(verbosity=0, sc_f=2, sc_l=1)
    with gpu.Handle(p, 160, 96, 1) as hd:
        out = hd.run_host(A[None], B[None])[0]
    want = port.run_dense(to_orc(p), A, B)
    assert np.isfinite(out).all()
    assert epe(out, want)[0] <= MEAN_TOL


def test_large_motion_outliers_and_borders(gpu, port):
    """Patches that start or move outside the valid box are reset (patch.cpp:133-139,197-206)."""
    A, B, _ = make_pair(120, 200, seed=9)
    B = np.roll(B, 37, axis=1)        # far beyond the search range at the fine levels
    p = gpu.operating_point(2, 200).copy(verbosity=0, sc_f=3, sc_l=1)
    with gpu.Handle(p, 200, 120, 1) as hd:
        out = hd.run_host(A[None], B[None])[0]
    want = port.run_dense(to_orc(p), A, B)
    m, mx = epe(out, want)
    assert np.isfinite(out).all() and m <= MEAN_TOL, (m, mx)


def test_deterministic(gpu):
    A, B, _ = make_pair(436, 1024, seed=2)
    p = gpu.operating_point(2, 1024).copy(verbosity=0, usefbcon=1)
    with gpu.Handle(p, 1024, 436, 2) as hd:
        o1 = hd.run_host(np.stack([A, A]), np.stack([B, B])).copy()
        o2 = hd.run_host(np.stack([A, A]), np.stack([B, B]))
    assert np.array_equal(o1, o2) and np.array_equal(o1[0], o1[1])


# ------------------------------------------------------------------------------ full bench size, properties only
def test_full_batch_properties(gpu, port):
    """BASELINE.json configs[0] at the bench batch size: every copy of a pair gives the same bits,
    distinct pairs match the oracle, device-pointer and host-pointer entry points agree."""
    import torch
    w0, h0, nb = 1024, 436, 592    # > 2 x 148 pairs: takes the large-batch (one warp per pair) SOR schedule
    p = gpu.operating_point(2, w0).copy(verbosity=0)
    pairs = [make_pair(h0, w0, seed=s)[:2] for s in range(2)]
    A = np.stack([pairs[i % 2][0] for i in range(nb)])
    B = np.stack([pairs[i % 2][1] for i in range(nb)])
    with gpu.Handle(p, w0, h0, nb) as hd:
        dA, dB = torch.from_numpy(A).cuda(), torch.from_numpy(B).cuda()
        dOut = torch.empty((nb, h0, w0, 2), dtype=torch.float32, device="cuda")
        torch.cuda.synchronize()
        hd.run_device(dA.data_ptr(), dB.data_ptr(), nb, dOut.data_ptr())
        hd.sync()
        out = dOut.cpu().numpy()
        host = hd.run_host(A[:6], B[:6])
    for s in range(2):
        assert all(np.array_equal(out[s], out[i]) for i in range(s, nb, 2))
        assert epe(out[s], port.run_dense(to_orc(p), *pairs[s]))[0] <= MEAN_TOL
    assert np.array_equal(host, out[:6])


# ------------------------------------------------------------------------------ one pair tiled over several strips / GPUs
TILED_CASES = [
    # id, mode, ch, w0, h0, overrides (sc_f = 2, sc_l = 0: levels of h0 / 4, h0 / 2, h0 rows)
    ("of_int", 1, 1, 640, 512, {}),
    ("of_int_fb_l1", 1, 1, 640, 512, dict(usefbcon=1, costfct=1)),
    ("de_int", 2, 1, 640, 512, {}),
    ("of_rgb_fb", 1, 3, 320, 416, dict(usefbcon=1)),
    ("of_int_no_tv", 1, 1, 640, 512, dict(usetvref=0, usefbcon=1)),
    ("of_int_psz12_2sweeps", 1, 1, 500, 390, dict(patchsz=12, patove=0.75, tv_solverit=2, tv_innerit=2)),
    ("of_int_solverit1_odd", 1, 1, 333, 407, dict(tv_solverit=1)),
]


@pytest.mark.parametrize("cfg", TILED_CASES, ids=[c[0] for c in TILED_CASES])
def test_tiled_strips_on_one_gpu_are_bit_identical(gpu, cfg):
    """disb200_tiled_* with the same device named 2, 3 and 4 times: the strips, the patch ownership, the all-gathers, the
    halo rows and the cross-strip banded SOR are the code path of real multi-GPU tiling (tests/test_tiled_multi_gpu.py
    runs it on distinct devices); the result must equal the untiled handle bit for bit."""
    _, mode, ch, w0, h0, kw = cfg
    A, B, _ = make_pair(h0, w0, channels=ch, mode="OF" if mode == 1 else "DE", seed=6)
    p = to_dis(gpu, base_orc(mode, ch, sc_f=2, sc_l=0, **kw))
    with gpu.Handle(p, w0, h0, max_batch=1) as hd:
        want = hd.run_host(A[None], B[None])[0]
    for ndev in (2, 3, 4):
        with gpu.Tiled(p, w0, h0, [0] * ndev) as td:
            tiled0, rows0, gy0 = td.plan(0)
            assert tiled0 and rows0[0] == 0
            got = td.run(A, B)
            got2 = td.run(A, B)              # exchange rows are reused under a new epoch
            assert td.launch_count() > 0
        assert all(r % 32 == 0 for r in rows0[:-1]) and rows0 == sorted(rows0) and gy0 == sorted(gy0)
        assert np.array_equal(got, want), (ndev, float(np.abs(got - want).max()))
        assert np.array_equal(got2, want)


def test_tiled_small_frame_is_replicated(gpu):
    """Levels of up to 96 rows are not cut: every device computes them; the result is the untiled one."""
    A, B, _ = make_pair(120, 200, seed=2)
    p = gpu.operating_point(2, 200).copy(verbosity=0)
    with gpu.Handle(p, 200, 120, 1) as hd:
        want = hd.run_host(A[None], B[None])[0]
    with gpu.Tiled(p, 200, 120, [0, 0]) as td:
        assert not any(td.plan(l)[0] for l in range(p.sc_l, p.sc_f + 1))
        assert np.array_equal(td.run(A, B), want)


@pytest.mark.parametrize("mode,ch,kw", [(1, 1, {}), (2, 1, {}), (1, 3, dict(usefbcon=1))])
def test_antiphase_sub_batches_give_the_same_bits(gpu, mode, ch, kw):
    """DisTuning.antiphase: two sub-batches enqueued level by level side by side, their refinements alternating (one
    assembles its linear system while the other runs the one-warp SOR).  A schedule, not an algorithm: identical output."""
    w0, h0, n = 320, 200, 7
    pairs = [make_pair(h0, w0, channels=ch, mode="OF" if mode == 1 else "DE", seed=30 + s)[:2] for s in range(n)]
    A, B = np.stack([a for a, _ in pairs]), np.stack([b for _, b in pairs])
    p = gpu.operating_point(2, w0, mode, ch).copy(verbosity=0, sc_f=3, sc_l=1, **kw)
    with gpu.Handle(p, w0, h0, n) as hd:
        want = hd.run_host(A, B)
    try:
        gpu.set_tuning(streams=2, antiphase=1, tv_schedule=gpu.TV_SPLIT)
        with gpu.Handle(p, w0, h0, n) as hd:
            got = hd.run_host(A, B)
            got2 = hd.run_host(A[:3], B[:3])
    finally:
        gpu.set_tuning()
    assert np.array_equal(got, want) and np.array_equal(got2, want[:3])


@pytest.mark.parametrize("mode,kw", [(1, {}), (2, {}), (1, dict(usefbcon=1, costfct=1)), (1, dict(costfct=2, patnorm=0, max_iter=24, min_iter=2))])
def test_staged_tap_windows_give_the_same_bits(gpu, mode, kw):
    """The 8 x 8 grey inverse search reads its bilinear taps from a shared-memory copy of each patch's 9 x 9 window;
    DisTuning.patch_map = 1 reads them from global memory.  Same values in the same order: identical output, also where
    the groups of a warp need different iteration counts (min_iter < max_iter)."""
    w0, h0, n = 328, 204, 5
    pairs = [make_pair(h0, w0, channels=1, mode="OF" if mode == 1 else "DE", seed=50 + s)[:2] for s in range(n)]
    A, B = np.stack([a for a, _ in pairs]), np.stack([b for _, b in pairs])
    p = gpu.operating_point(2, w0, mode, 1).copy(verbosity=0, sc_f=3, sc_l=0, **kw)
    with gpu.Handle(p, w0, h0, n) as hd:
        want = hd.run_host(A, B)
    try:
        gpu.set_tuning(patch_map=1)
        with gpu.Handle(p, w0, h0, n) as hd:
            got = hd.run_host(A, B)
    finally:
        gpu.set_tuning()
    assert np.array_equal(got, want) and np.abs(want).max() > 0.5
