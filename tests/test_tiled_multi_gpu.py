"""One frame pair tiled spatially over SEVERAL PHYSICAL GPUs (NVLink P2P): run with `gpurun --gpus N -- python -m pytest
tests/test_tiled_multi_gpu.py -m gpu`.  On a one-GPU box these tests skip; the same code path is covered there with
repeated device ids (tests/test_parity_gpu.py::test_tiled_strips_on_one_gpu_are_bit_identical)."""
import numpy as np
import pytest

from of_dis_b200.synth import make_pair
from tests.helpers import base_orc, to_dis

pytestmark = pytest.mark.gpu


def _need(gpu, n):
    if gpu.device_count() < n:
        pytest.skip("needs %d GPUs, this box has %d" % (n, gpu.device_count()))


@pytest.mark.parametrize("ndev", [2, 4, 8])
@pytest.mark.parametrize("kw", [dict(), dict(usefbcon=1, costfct=1), dict(mode=2)], ids=["of", "of_fb_l1", "de"])
def test_tiled_over_physical_gpus_is_bit_identical(gpu, ndev, kw):
    _need(gpu, ndev)
    kw = dict(kw)
    mode = kw.pop("mode", 1)
    w0, h0 = 960, 1088
    A, B, _ = make_pair(h0, w0, mode="OF" if mode == 1 else "DE", seed=8)
    p = to_dis(gpu, base_orc(mode, 1, sc_f=3, sc_l=0, **kw))
    with gpu.Handle(p, w0, h0, max_batch=1) as hd:
        want = hd.run_host(A[None], B[None])[0]
    with gpu.Tiled(p, w0, h0, list(range(ndev))) as td:
        assert td.plan(0)[0]
        got = td.run(A, B)
        got2 = td.run(A, B)
    assert np.array_equal(got, want), float(np.abs(got - want).max())
    assert np.array_equal(got2, want)


@pytest.mark.parametrize("ndev", [2, 4])
def test_tiled_4k_heavy_and_light_lines(gpu, ndev):
    """BASELINE configs[3]: one 3840x2160 pair, forward-backward consistency + L1, both C4 parameter lines of SURVEY 8(d)."""
    _need(gpu, ndev)
    w0, h0 = 3840, 2160
    A, B, _ = make_pair(h0, w0, seed=0)
    for line in (dict(sc_f=7, sc_l=2, max_iter=16, min_iter=16, patchsz=8, patove=0.4, usefbcon=1, costfct=1),
                 dict(sc_f=7, sc_l=1, max_iter=16, min_iter=16, patchsz=12, patove=0.75, usefbcon=1, costfct=1)):
        p = gpu.operating_point(2, w0).copy(verbosity=0, **line)
        with gpu.Handle(p, w0, h0, max_batch=1) as hd:
            want = hd.run_host(A[None], B[None])[0]
        with gpu.Tiled(p, w0, h0, list(range(ndev))) as td:
            got = td.run(A, B)
        assert np.array_equal(got, want), (line, float(np.abs(got - want).max()))
