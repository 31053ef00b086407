"""CPU suite (-m "not gpu"): the oracle against its golden vectors, host logic, C-ABI exports."""
import ctypes
import importlib.util
import os
import re
import struct
import subprocess
import sys

import numpy as np
import pytest

from oracle import bindings as ob
from tests.helpers import TINY_CASES, base_orc, epe

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _ref_cases():
    spec = importlib.util.spec_from_file_location("make_golden", os.path.join(ROOT, "tests", "golden", "make_golden.py"))
    m = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(m)
    return m.REF_CASES


REF_CASES = _ref_cases()


# ---------------------------------------------------------------- OpenCV steps vs python-cv2 golden
@pytest.mark.parametrize("tag", ["g", "c"])
def test_pyramid_steps_match_cv2(port, cv2_golden, tag):
    g = cv2_golden
    lv = g[tag + "_l0"]
    for l in (1, 2, 3):
        lv = port.halve(lv)
        assert np.array_equal(lv, g["%s_l%d" % (tag, l)]), "2x2 box mean differs from cv2.resize at level %d" % l
        dx, dy = port.sobel(lv)
        assert np.array_equal(dx, g["%s_dx%d" % (tag, l)])
        assert np.array_equal(dy, g["%s_dy%d" % (tag, l)])


@pytest.mark.parametrize("tag", ["g", "c"])
def test_pyramid_level_from_u8_matches_cv2(port, cv2_golden, tag):
    """a1 + a2 end to end: u8 frame -> replicate pad to /16 -> level 2 -> Sobel -> border pads."""
    g = cv2_golden
    u8 = g[tag + "_u8"]
    img, dx, dy = port.pyramid_level(u8, sc_f=4, level=2, pad=8)
    want_img = g[tag + "_pad2_rep"].reshape(img.shape)
    want_dx = g[tag + "_pad2_zero"].reshape(dx.shape)
    assert np.array_equal(img, want_img)
    assert np.array_equal(dx, want_dx)
    assert port.pad_dims(72, 44, 4) == (80, 48, 4, 2)


@pytest.mark.parametrize("nop", [1, 2])
@pytest.mark.parametrize("sc_l", [1, 2, 3])
def test_upsample_crop_matches_cv2(port, cv2_golden, nop, sc_l):
    fl = cv2_golden["up_in_%d_%d" % (nop, sc_l)]
    want = cv2_golden["up_out_%d_%d" % (nop, sc_l)]
    h0, w0 = want.shape[:2]
    got = port.upsample_crop(fl, sc_l, 1, 2, w0, h0)
    assert np.abs(got - want).max() <= 1e-5   # cv2 differs by fp32 rounding at most


def test_upsample_sc_l_zero_is_a_crop(port):
    fl = np.arange(6 * 8 * 2, dtype=np.float32).reshape(6, 8, 2)
    got = port.upsample_crop(fl, 0, 1, 2, 5, 3)
    assert np.array_equal(got, fl[2:5, 1:6])


# ---------------------------------------------------------------- oracle vs the reference's own code
@pytest.mark.parametrize("name", sorted(REF_CASES))
def test_port_matches_reference_golden(port, ref_golden, name):
    mode, noc, h0, w0, kw = REF_CASES[name]
    A, B = ref_golden[name + "_A"], ref_golden[name + "_B"]
    p = base_orc(mode, noc, **kw)
    W, H, _, _ = port.pad_dims(w0, h0, p.sc_f)
    PA = port.pyramid(A, p.sc_f, p.sc_l, p.psz)
    PB = port.pyramid(B, p.sc_f, p.sc_l, p.psz)
    got = port.ofclass(p, PA, PB, W, H)
    want = ref_golden[name + "_low"]
    assert np.array_equal(got, want), "restatement differs from the reference build: EPE mean %g max %g" % epe(got, want)


@pytest.mark.parametrize("mode,noc", [(1, 1), (1, 3), (2, 1), (2, 3)])
def test_port_matches_reference_live(port, mode, noc):
    if not ob.have_ref(mode, noc):
        pytest.skip("oracle/_ref not built on this machine")
    from of_dis_b200.synth import make_pair
    ref = ob.Ref(mode, noc)
    A, B, _ = make_pair(70, 110, channels=noc, mode="OF" if mode == 1 else "DE", seed=5)
    for kw in (dict(), dict(usefbcon=1), dict(costfct=2, min_iter=3, max_iter=20, res_thresh=.3), dict(sc_f=2, sc_l=0, usetvref=0)):
        p = base_orc(mode, noc, **kw)
        W, H, _, _ = port.pad_dims(110, 70, p.sc_f)
        PA = port.pyramid(A, p.sc_f, p.sc_l, p.psz)
        PB = port.pyramid(B, p.sc_f, p.sc_l, p.psz)
        assert np.array_equal(port.ofclass(p, PA, PB, W, H), ref.ofclass(p, PA, PB, W, H)), kw
        # per-stage: patch state and densified flow on the finest level, refinement alone
        wl, hl = W >> p.sc_l, H >> p.sc_l
        a = (PA[0][p.sc_l], PA[1][p.sc_l], PA[2][p.sc_l])
        b = (PB[0][p.sc_l], PB[1][p.sc_l], PB[2][p.sc_l])
        ro, rr = port.patchgrid_level(p, a, b, wl, hl, p.sc_l), ref.patchgrid_level(p, a, b, wl, hl, p.sc_l)
        for k in ("p_iter_fw", "pt_iter_fw", "pweight_fw", "cnt_fw", "dense_fw"):
            assert np.array_equal(ro[k], rr[k]), (kw, k)
        fl = ro["dense_fw"]
        assert np.array_equal(port.varref_level(p, a[0], b[0], wl, hl, p.sc_l, 0, fl),
                              ref.varref_level(p, a[0], b[0], wl, hl, p.sc_l, 0, fl)), kw


@pytest.mark.parametrize("mode,noc,psz,ov", [(1, 1, 5, .4), (1, 1, 9, .5), (1, 3, 5, .4), (1, 1, 6, .4), (2, 1, 7, .4), (1, 1, 2, 0.), (1, 1, 3, 0.),
                                             (1, 3, 7, .6), (2, 3, 15, .7), (1, 1, 10, .6)])
def test_port_matches_reference_on_every_patch_size(port, mode, noc, psz, ov):
    """The 20-parameter interface takes any patch size (oflow.cpp:90-91).  Odd sizes sample (psz - 1)^2 positions into
    vectors of psz^2 entries whose tail LossComputeErrorImage partly skips (patch.cpp:226-232,316-317,377-378); the
    restatement must follow the reference through all of that, bit for bit (oracle/_ref zero-initialises the vectors)."""
    if not ob.have_ref(mode, noc):
        pytest.skip("oracle/_ref not built on this machine")
    from of_dis_b200.synth import make_pair
    ref = ob.Ref(mode, noc)
    A, B, _ = make_pair(96, 160, channels=noc, mode="OF" if mode == 1 else "DE", seed=1)
    for kw in (dict(), dict(usefbcon=1, costfct=1)):
        p = base_orc(mode, noc, psz=psz, patove=ov, sc_f=2, sc_l=0, **kw)
        W, H, _, _ = port.pad_dims(160, 96, p.sc_f)
        PA = port.pyramid(A, p.sc_f, p.sc_l, p.psz)
        PB = port.pyramid(B, p.sc_f, p.sc_l, p.psz)
        assert np.array_equal(port.ofclass(p, PA, PB, W, H), ref.ofclass(p, PA, PB, W, H)), kw


@pytest.mark.parametrize("case", TINY_CASES)
def test_port_matches_reference_on_tiny_levels(port, case):
    """Pyramid levels down to one pixel: the restatement against oracle/_ref.  (A level one ROW high with the refinement
    on is outside the domain: the reference itself dies with a segmentation fault there.)"""
    mode, noc, w0, h0, sc_f, sc_l, kw = case
    if not ob.have_ref(mode, noc):
        pytest.skip("oracle/_ref not built on this machine")
    from of_dis_b200.synth import make_pair
    A, B, _ = make_pair(h0, w0, channels=noc, mode="OF" if mode == 1 else "DE", seed=1)
    p = base_orc(mode, noc, sc_f=sc_f, sc_l=sc_l, **kw)
    W, H, _, _ = port.pad_dims(w0, h0, p.sc_f)
    PA = port.pyramid(A, p.sc_f, p.sc_l, p.psz)
    PB = port.pyramid(B, p.sc_f, p.sc_l, p.psz)
    want = ob.Ref(mode, noc).ofclass(p, PA, PB, W, H)
    assert np.isfinite(want).all()
    assert np.array_equal(port.ofclass(p, PA, PB, W, H), want)


# ---------------------------------------------------------------- SELECTCHANNEL == 2 (gradient-magnitude input)
def test_gradmag_pyramid_steps_match_cv2(port, gradmag_golden):
    """run_dense.cpp:139-148 on arbitrary floats: level 0 and the .5 cascade (OpenCV's own summation order, including
    the w % 4 tail columns: levels of 40, 20 and 10 columns) bit for bit; the IPP dispatch of the same wheel is
    within one ulp of it."""
    g = gradmag_golden
    f = np.pad(g["g_u8"], ((2, 2), (4, 4)), mode="edge").astype(np.float32)
    lv = port.gradient_magnitude(f)
    assert np.array_equal(lv, g["g_l0"])
    for l in (1, 2, 3):
        lv = port.halve(lv)
        dx, dy = port.sobel(lv)
        assert np.array_equal(lv, g["g_l%d" % l]) and np.array_equal(dx, g["g_dx%d" % l]) and np.array_equal(dy, g["g_dy%d" % l]), l
        assert np.abs(g["g_ipp_l%d" % l] - lv).max() <= np.spacing(np.float32(lv.max()))
        I, X, Y = port.pyramid_level(g["g_u8"], 4, l, 8, gradmag=True)
        assert np.array_equal(I[8:-8, 8:-8, 0], g["g_l%d" % l]) and np.array_equal(X[8:-8, 8:-8, 0], g["g_dx%d" % l])
        assert np.array_equal(Y[8:-8, 8:-8, 0], g["g_dy%d" % l])
    # the intensity pyramid is not touched by the switch being off again
    assert np.array_equal(port.pyramid_level(g["g_u8"], 4, 0, 8)[0][8:-8, 8:-8, 0], f)


@pytest.mark.parametrize("mode,name", [(1, "of_grad_op2like"), (2, "de_grad_op2like")])
def test_port_matches_reference_golden_gradmag(port, gradmag_golden, mode, name):
    """Flow of the reference's own sources compiled with -DSELECTCHANNEL=2 on the gradient-magnitude pyramid."""
    g = gradmag_golden
    A, B = g[name + "_A"], g[name + "_B"]
    p = base_orc(mode, 1, gradmag=1)
    W, H, _, _ = port.pad_dims(150, 90, p.sc_f)
    PA, PB = port.pyramid(A, p.sc_f, p.sc_l, p.psz, gradmag=True), port.pyramid(B, p.sc_f, p.sc_l, p.psz, gradmag=True)
    assert np.array_equal(port.ofclass(p, PA, PB, W, H), g[name + "_low"])
    if ob.have_ref(mode, 1, gradmag=True):
        assert np.array_equal(ob.Ref(mode, 1, gradmag=True).ofclass(p, PA, PB, W, H), g[name + "_low"])
    # the driver entry point builds the same pyramid from the parameter block
    out, low = port.run_dense(p, A, B, want_lowres=True)
    assert np.array_equal(low, g[name + "_low"])
    assert not np.array_equal(low, port.run_dense(p.copy(gradmag=0), A, B, want_lowres=True)[1])
    gt = g[name + "_gt"]
    assert epe(out, gt)[0] < 1.5      # a flow, if a worse one than on intensities (texture edges only)


def test_port_flow_is_sane(port):
    """EPE against the synthetic ground truth (not a parity check, a sanity check of the generator + oracle)."""
    from of_dis_b200.synth import make_pair
    A, B, gt = make_pair(218, 512, channels=1, mode="OF", seed=2)
    p = port.operating_point(2, 512, 1, 1).copy(verbosity=0)
    out = port.run_dense(p, A, B)
    assert out.shape == (218, 512, 2)
    assert epe(out, gt)[0] < 0.8


def test_identical_frames_give_zero_flow(port):
    from of_dis_b200.synth import make_pair
    A, _, _ = make_pair(100, 160, channels=1, mode="OF", seed=1)
    p = base_orc(1, 1)
    out = port.run_dense(p, A, A)
    assert np.all(out == 0)


# ---------------------------------------------------------------- host logic of the product (no GPU needed)
def test_operating_points_match_oracle(dis, port):
    for op in (1, 2, 3, 4, 7):
        for width in (320, 640, 1024, 1280, 1920, 3840):
            for mode, ch in ((1, 1), (1, 3), (2, 1)):
                p = dis.operating_point(op, width, mode, ch)
                q = port.operating_point(op, width, mode, ch)
                for f, _ in ob.OrcParams._fields_:
                    assert getattr(p, {"psz": "patchsz", "noc": "channels"}.get(f, f)) == getattr(q, f), (op, width, f)
    p = dis.operating_point(2, 1024)
    # run_dense.cpp:260-266 + README: op2 on a 1024 wide frame = 5 3 12 12 .05 .95 0 8 .4 0 1 0 1 10 10 5 1 3 1.6
    assert p.cli_args()[:19] == pytest.approx([5, 3, 12, 12, .05, .95, 0, 8, .4, 0, 1, 0, 1, 10, 10, 5, 1, 3, 1.6])


def test_padded_dims_match_oracle(dis, port):
    for (w0, h0, sc_f) in ((1024, 436, 5), (1280, 720, 6), (3840, 2160, 7), (1920, 1080, 6), (33, 17, 3), (64, 64, 4)):
        assert dis.padded_dims(w0, h0, sc_f) == port.pad_dims(w0, h0, sc_f)
    assert dis.padded_dims(1024, 436, 5) == (1024, 448, 0, 6)


def test_cabi_exports_every_declared_symbol(dis):
    hdr = open(os.path.join(ROOT, "include", "disb200.h")).read()
    names = sorted(set(re.findall(r"\b(disb200_[a-z0-9_]+)\s*\(", hdr)))
    assert len(names) >= 20
    L = dis.lib()
    for n in names:
        assert hasattr(L, n), "libdisb200.so does not export " + n
    assert b"sm_100a" in L.disb200_version()


def test_struct_layout_matches_header(dis):
    # 20 four-byte fields in CLI order + mode, channels, sor_mode, gradmag
    assert ctypes.sizeof(dis.DisParams) == 24 * 4
    hdr = open(os.path.join(ROOT, "include", "disb200.h")).read()
    body = hdr[hdr.index("typedef struct DisParams {"):hdr.index("} DisParams;")]
    fields = re.findall(r"^\s*(?:int|float)\s+([a-z_]+);", body, re.M)
    assert fields == [f for f, _ in dis.DisParams._fields_]


def test_compute_entry_points_fail_loudly_without_gpu(dis):
    if dis.device_count() > 0:
        pytest.skip("a GPU is present")
    p = dis.operating_point(2, 256)
    with pytest.raises(dis.DisError) as e:
        dis.Handle(p, 256, 128, 1)
    assert e.value.code == 2 and "no CPU path" in str(e.value)


def test_clis_fail_loudly_without_gpu(dis, tmp_path):
    """The run_* binaries (all six SELECTMODE x SELECTCHANNEL builds) have no CPU path either: usage on bad argv,
    a clear error and a non-zero exit code when no device is present."""
    exe_dir = os.path.join(ROOT, "of_dis_b200", "bin")
    names = ["run_OF_INT", "run_OF_RGB", "run_DE_INT", "run_DE_RGB", "run_OF_GRAD", "run_DE_GRAD"]
    for n in names + ["disb200_batch"]:
        assert os.access(os.path.join(exe_dir, n), os.X_OK), n
    r = subprocess.run([os.path.join(exe_dir, "run_OF_INT")], capture_output=True, text=True)
    assert r.returncode != 0 and "usage" in (r.stderr + r.stdout)
    if dis.device_count() > 0:
        pytest.skip("a GPU is present")
    img = np.random.default_rng(0).integers(0, 256, (60, 100), dtype=np.uint8)
    for name, data, hdr in (("a.pgm", img, b"P5\n100 60\n255\n"), ("a.ppm", np.repeat(img[..., None], 3, -1), b"P6\n100 60\n255\n")):
        with open(tmp_path / name, "wb") as f:
            f.write(hdr + data.tobytes())
    for n in names:
        src = str(tmp_path / ("a.ppm" if n.endswith("RGB") else "a.pgm"))
        r = subprocess.run([os.path.join(exe_dir, n), src, src, str(tmp_path / "o.flo")], capture_output=True, text=True)
        assert r.returncode == 2 and "no CPU path" in (r.stderr + r.stdout), (n, r.stdout, r.stderr)
        assert not os.path.exists(tmp_path / "o.flo")


def test_invalid_parameters_are_rejected(dis):
    p = dis.operating_point(2, 256)
    for kw in (dict(mode=3), dict(channels=2), dict(sc_l=4, sc_f=3), dict(patove=1.0), dict(patchsz=1), dict(patchsz=17), dict(costfct=3), dict(tv_solverit=-1), dict(sor_mode=2), dict(gradmag=2),
               dict(gradmag=1, channels=3)):
        with pytest.raises(dis.DisError) as e:
            dis.Handle(p.copy(**kw), 256, 128, 1)
        assert e.value.code in (1, 4), kw


def test_flo_and_pfm_writers_use_the_reference_framing(dis, tmp_path):
    fl = np.arange(3 * 5 * 2, dtype=np.float32).reshape(3, 5, 2) - 7
    f = str(tmp_path / "a.flo")
    dis.save_flo(f, fl)
    raw = open(f, "rb").read()
    assert raw[:4] == b"PIEH" and struct.unpack("<ii", raw[4:12]) == (5, 3)          # run_dense.cpp:29-32
    assert np.array_equal(np.frombuffer(raw[12:], np.float32).reshape(3, 5, 2), fl)
    # SaveFlowFile writes img.channels() floats per pixel: depth (1) and scene flow (4) share the framing (run_dense.cpp:37-51)
    for nc in (1, 4):
        x = (np.arange(3 * 5 * nc, dtype=np.float32).reshape(3, 5, nc) * 0.5 - 3)
        f = str(tmp_path / ("c%d.flo" % nc))
        dis.save_flo(f, x[..., 0] if nc == 1 else x)
        raw = open(f, "rb").read()
        assert raw[:4] == b"PIEH" and struct.unpack("<ii", raw[4:12]) == (5, 3) and len(raw) == 12 + 3 * 5 * nc * 4
        assert np.array_equal(np.frombuffer(raw[12:], np.float32).reshape(3, 5, nc), x)
    with pytest.raises(dis.DisError):
        dis.save_flo(str(tmp_path / "bad.flo"), np.zeros((3, 5, 3), np.float32))
    d = fl[..., 0]
    f = str(tmp_path / "a.pfm")
    dis.save_pfm(f, d)
    raw = open(f, "rb").read()
    head = b"Pf\n5 3\n-1.000000\n"                                                     # run_dense.cpp:70
    assert raw.startswith(head)
    body = np.frombuffer(raw[len(head):], np.float32).reshape(3, 5)
    assert np.array_equal(body, -d[::-1])                                              # bottom-up, negated (:72-79)


def test_flo_reader_round_trip_and_error_paths(dis, tmp_path):
    """ReadFlowFile (run_dense.cpp:84-128): what SaveFlowFile wrote comes back bit for bit for depth, flow and scene
    flow; where the reference only prints (short / long file) the library returns DISB200_EIO."""
    rng = np.random.default_rng(3)
    for nc in (1, 2, 4):
        x = rng.standard_normal((7, 11, nc)).astype(np.float32)
        f = str(tmp_path / ("r%d.flo" % nc))
        dis.save_flo(f, x[..., 0] if nc == 1 else x)
        got = dis.read_flo(f, nc)
        assert got.shape == ((7, 11) if nc == 1 else (7, 11, nc)) and np.array_equal(got.reshape(7, 11, nc), x)
    f2 = str(tmp_path / "r2.flo")
    raw = open(f2, "rb").read()
    with pytest.raises(dis.DisError) as e:          # a flow file read as scene flow is too short
        dis.read_flo(f2, 4)
    assert e.value.code == 5
    with pytest.raises(dis.DisError):               # ... and too long when read as depth
        dis.read_flo(f2, 1)
    open(tmp_path / "trunc.flo", "wb").write(raw[:-4])
    open(tmp_path / "long.flo", "wb").write(raw + b"\0")
    open(tmp_path / "tag.flo", "wb").write(b"XXXX" + raw[4:])
    open(tmp_path / "hdr.flo", "wb").write(raw[:10])
    for name in ("trunc.flo", "long.flo", "tag.flo", "hdr.flo", "missing.flo"):
        with pytest.raises(dis.DisError) as e:
            dis.read_flo(str(tmp_path / name), 2)
        assert e.value.code == 5, name
    with pytest.raises(dis.DisError) as e:
        dis.read_flo(f2, 3)
    assert e.value.code == 1


def test_pnm_reader(dis, tmp_path):
    rng = np.random.default_rng(0)
    g = rng.integers(0, 256, (7, 9), dtype=np.uint8)
    c = rng.integers(0, 256, (7, 9, 3), dtype=np.uint8)   # RGB on disk
    pg, pc = str(tmp_path / "g.pgm"), str(tmp_path / "c.ppm")
    open(pg, "wb").write(b"P5\n# comment\n9 7\n255\n" + g.tobytes())
    open(pc, "wb").write(b"P6 9 7 255\n" + c.tobytes())
    assert np.array_equal(dis.read_pnm(pg, 1), g)
    assert np.array_equal(dis.read_pnm(pc, 3), c[..., ::-1])                          # BGR like cv::imread
    assert np.array_equal(dis.read_pnm(pg, 3), np.repeat(g[..., None], 3, -1))
    grey = dis.read_pnm(pc, 1).astype(int)
    want = (c[..., 0].astype(int) * 4899 + c[..., 1].astype(int) * 9617 + c[..., 2].astype(int) * 1868 + 8192) >> 14
    assert np.array_equal(grey, want)


# ---------------------------------------------------------------- multi-GPU host logic on CPU (gloo, world 2)
def test_shard_range_covers_everything():
    from of_dis_b200.sharding import frames_for_pairs, shard_range
    for n in (0, 1, 7, 256, 257):
        for world in (1, 2, 3, 8):
            blocks = [shard_range(n, r, world) for r in range(world)]
            assert blocks[0][0] == 0 and blocks[-1][1] == n
            assert all(blocks[i][1] == blocks[i + 1][0] for i in range(world - 1))
            assert max(b[1] - b[0] for b in blocks) - min(b[1] - b[0] for b in blocks) <= 1
    assert frames_for_pairs(32, 64) == (32, 65)


_GLOO_WORKER = r"""
import os, sys
sys.path.insert(0, sys.argv[1])
import numpy as np, torch, torch.distributed as dist
from of_dis_b200.sharding import shard_range
from of_dis_b200.synth import make_pair
from oracle import bindings as ob
dist.init_process_group("gloo")
rank, world = dist.get_rank(), dist.get_world_size()
port = ob.Port()
n = 5
lo, hi = shard_range(n, rank, world)
p = port.operating_point(2, 160, 1, 1).copy(verbosity=0)
sums = torch.zeros(n, dtype=torch.float64)
for i in range(lo, hi):                      # each rank processes only its own block: no data-path collective
    A, B, _ = make_pair(64, 160, seed=i)
    sums[i] = float(np.abs(port.run_dense(p, A, B)).sum())
cnt = torch.tensor([hi - lo])
dist.all_reduce(sums); dist.all_reduce(cnt)   # bookkeeping only (what bench.py does with the timings)
if rank == 0:
    full = [float(np.abs(port.run_dense(p, *make_pair(64, 160, seed=i)[:2])).sum()) for i in range(n)]
    assert int(cnt) == n and np.allclose(sums.numpy(), full, rtol=0, atol=0), (sums, full)
    print("OK")
dist.destroy_process_group()
"""


def test_pair_sharding_world2_gloo(tmp_path, port):
    script = tmp_path / "w.py"
    script.write_text(_GLOO_WORKER)
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
                        "--master-addr", "127.0.0.1", "--master-port", "29741", str(script), ROOT],
                       capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and "OK" in r.stdout, r.stdout[-2000:] + r.stderr[-2000:]


def test_oracle_redblack_order_is_a_different_iterate_of_the_same_problem(port):
    """orc_set_sor_order(1) restates the product's DISB200_SOR_REDBLACK schedule (not in the reference)."""
    from of_dis_b200.synth import make_pair
    from tests.helpers import base_orc
    A, B, _ = make_pair(100, 160, seed=2)
    q = base_orc(1, 1, sc_f=1, sc_l=1)
    a, b = port.pyramid_level(A, 1, 1, 8)[0], port.pyramid_level(B, 1, 1, 8)[0]
    fl = np.full((50, 80, 2), 0.25, np.float32)
    lex = port.varref_level(q, a, b, 80, 50, 1, 0, fl)
    try:
        port.set_sor_order(1)
        rb = port.varref_level(q, a, b, 80, 50, 1, 0, fl)
        rb_many = port.varref_level(q.copy(tv_solverit=200), a, b, 80, 50, 1, 0, fl)
    finally:
        port.set_sor_order(0)
    lex_many = port.varref_level(q.copy(tv_solverit=200), a, b, 80, 50, 1, 0, fl)
    assert np.array_equal(lex, port.varref_level(q, a, b, 80, 50, 1, 0, fl))      # order was reset
    assert not np.array_equal(rb, lex) and np.abs(rb - lex).mean() < 0.1
    assert np.abs(rb_many - lex_many).max() < 5e-2                                  # both orders converge to one solution


def test_oracle_ncc_definition(port):
    """costfct 10 is defined by this build (no reference branch): deterministic, sane flow, insensitive to gain."""
    from of_dis_b200.synth import make_pair
    A, B, gt = make_pair(120, 256, seed=1)
    q = port.operating_point(2, 256).copy(verbosity=0, costfct=10, usetvref=0)
    f1, f2 = port.run_dense(q, A, B), port.run_dense(q, A, B)
    assert np.array_equal(f1, f2)
    e = np.sqrt(((f1 - gt) ** 2).sum(-1)).mean()
    B2 = np.clip(B.astype(np.float32) * 0.7 + 20, 0, 255).astype(np.uint8)
    e2 = np.sqrt(((port.run_dense(q, A, B2) - gt) ** 2).sum(-1)).mean()
    e2_l2 = np.sqrt(((port.run_dense(q.copy(costfct=0), A, B2) - gt) ** 2).sum(-1)).mean()
    assert e < 1.5 and e2 < e + 0.1 and e2 < e2_l2


# ---------------------------------------------------------------- bench.py, reference arm (runs on host cores only)
def test_bench_reference_arm_prints_the_contract_line():
    """`bench.py --impl reference` times the reference's CPU code (oracle/_ref, else the restatement) and needs no GPU:
    one JSON line with the contract's keys, the same metric / unit / config as the GPU arm."""
    import json
    import subprocess
    import sys
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "2", "--warmup", "1", "--ref-seconds", "12"],
                       capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [l for l in r.stdout.splitlines() if l.startswith("{")]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["metric"] == "frame_pairs_per_s" and d["unit"] == "pairs/s"
    assert d["value"] > 0 and d["higher_is_better"] is True and d["n_gpus"] == 1 and d["steps"] == 2 and d["warmup"] == 1
    assert d["config"]["workload"].startswith("C1: run_OF_INT op2 1024x436 grey")
    assert d["cpu_baseline"]["kind"] in ("reference", "port") and d["cpu_baseline"]["cores"] >= 1
    assert d["cpu_baseline"]["value"] == d["value"] == d["e2e"]["value"]
    assert d["e2e"]["h2d_bytes_per_step"] == 0 and d["e2e"]["d2h_bytes_per_step"] == 0
    # core (the OFClass constructor) and pre/post (the OpenCV steps) separately, one thread and all cores
    cb = d["cpu_baseline"]
    assert cb["kind_prepost"] in ("cv2", "port") and cb["core_ms_1t"] > 0 and cb["prepost_ms_1t"] > 0
    assert cb["core_pairs_per_s"] >= cb["core_plus_prepost_pairs_per_s"] == cb["value"]
    # the reference arm never loads the product library
    assert d["native_so_loaded"] and not any("libdisb200" in x for x in d["native_so_loaded"]), d["native_so_loaded"]


def test_bench_parameter_overrides_are_for_kernel_experiments_only():
    """`bench.py --set k=v` changes the GPU arm's parameters only; it is refused unless the legs that would compare
    against (or time) the unmodified reference parameters are switched off, so no contract line can carry it by accident."""
    import subprocess
    import sys
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--set", "max_iter=3"], capture_output=True, text=True, timeout=300)
    assert r.returncode == 2 and "--set is for kernel experiments" in r.stderr, (r.returncode, r.stderr[-500:])
    assert not [l for l in r.stdout.splitlines() if l.startswith("{")]


@pytest.mark.parametrize("mode,noc,w0,h0", [(1, 1, 333, 217), (1, 3, 256, 150), (2, 1, 320, 200)])
def test_cv2_prepost_of_the_cpu_baseline_is_the_oracles(port, mode, noc, w0, h0):
    """bench.py times python-cv2 for the reference's OpenCV steps (oracle/cv2_prepost.py); what it computes must be what
    the parity oracle computes: pyramids bit for bit on u8 frames, the final resize within 1e-5 px."""
    from oracle import cv2_prepost as cp
    if not cp.HAVE_CV2:
        pytest.skip("python-cv2 not importable")
    from of_dis_b200.synth import make_pair
    A, _, _ = make_pair(h0, w0, channels=noc, mode="OF" if mode == 1 else "DE", seed=5)
    sc_f, sc_l, pad = 3, 1, 8
    PA = port.pyramid(A, sc_f, sc_l, pad)
    Ap, padl, padt = cp.pad_frame(A, sc_f)
    CA = cp.pyramid(Ap, sc_f, pad)
    W, H, pl, pt = port.pad_dims(w0, h0, sc_f)
    assert (padl, padt) == (pl, pt) and Ap.shape[:2] == (H, W)
    for k in range(3):
        for l in range(sc_l, sc_f + 1):
            assert np.array_equal(PA[k][l].reshape(CA[k][l].shape), CA[k][l]), (k, l)
    nop = 2 if mode == 1 else 1
    low = np.random.default_rng(0).standard_normal((H >> sc_l, W >> sc_l, nop)).astype(np.float32) * 5
    a = port.upsample_crop(low, sc_l, pl, pt, w0, h0)
    b = cp.upsample_crop(low, sc_l, padl, padt, w0, h0)
    assert a.shape == b.shape and np.abs(a - b).max() <= 1e-5
