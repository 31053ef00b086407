"""disb200_aggregate_dense_static: the allocation-free variant of the AggregateFlowDense cut (include/disb200.h).

No counterpart in the reference - an entry point for callers that cannot call cudaMalloc.  Its scratch (the CSR binning
of the forward-backward merge) is a static device array that every call shares and that keeps the previous call's
contents, so the test makes TWO calls in a row on different levels - the second one smaller, with a different
distribution of end positions, inside memory the first one filled - and holds both to the existing entry point's
output, bit for bit (same kernels, same order).  Collected last on purpose (pytest -x): a failure here hides nothing.
"""
import numpy as np
import pytest

from tests.helpers import base_orc, to_dis

pytestmark = pytest.mark.gpu


def _inputs(gpu, p, w_l, h_l, noc, psz, spread, seed):
    n = gpu.patch_count(p, w_l, h_l)
    nop, nv = p.nop, noc * psz * psz
    rng = np.random.default_rng(seed)
    p_iter = rng.standard_normal((n, nop)).astype(np.float32) * 3
    pweight = (rng.random((n, nv)) * 9).astype(np.float32)
    cg_p = rng.standard_normal((n, nop)).astype(np.float32) * 3
    cg_w = (rng.random((n, nv)) * 9).astype(np.float32)
    if spread == "regular":
        cg_pt = np.stack([rng.random(n) * (w_l + 20) - 10, rng.random(n) * (h_l + 20) - 10], -1).astype(np.float32)
    else:   # every complementary patch ends inside one cell
        cg_pt = np.stack([60 + rng.random(n) * 5, 40 + rng.random(n) * 5], -1).astype(np.float32)
    return p_iter, pweight, (cg_p, cg_pt, cg_w)


def _static(gpu, torch, p, w_l, h_l, p_iter, pweight, cg):
    """through the C-ABI with caller-owned device arrays (torch only holds the memory)"""
    dev = [torch.from_numpy(x).cuda() for x in (p_iter, pweight) + (tuple(cg) if cg else ())]
    out = torch.full((h_l, w_l, p.nop), float("nan"), dtype=torch.float32, device="cuda")
    torch.cuda.synchronize()
    gpu.aggregate_dense_static(p, w_l, h_l, dev[0].data_ptr(), dev[1].data_ptr(), [d.data_ptr() for d in dev[2:]] if cg else None,
                               out.data_ptr(), 0)
    torch.cuda.synchronize()
    return out.cpu().numpy()


@pytest.mark.parametrize("mode,noc,psz,ov", [(1, 1, 8, .4), (2, 3, 12, .75)])
def test_static_scratch_entry_matches_the_allocating_one_on_two_calls_in_a_row(gpu, mode, noc, psz, ov):
    torch = pytest.importorskip("torch")
    p = to_dis(gpu, base_orc(mode, noc, psz=psz, patove=ov))
    calls = [(412, 260, "collapsed", 3), (150, 90, "regular", 4), (150, 90, "collapsed", 5)]
    for w_l, h_l, spread, seed in calls:
        p_iter, pweight, cg = _inputs(gpu, p, w_l, h_l, noc, psz, spread, seed)
        want = gpu.aggregate_dense(p, w_l, h_l, p_iter, pweight, cg)
        got = _static(gpu, torch, p, w_l, h_l, p_iter, pweight, cg)
        assert np.isfinite(got).all()
        assert np.array_equal(got, want), (w_l, h_l, spread, float(np.abs(got - want).max()))
    # without the complementary grid no scratch is touched
    want = gpu.aggregate_dense(p, w_l, h_l, p_iter, pweight, None)
    assert np.array_equal(_static(gpu, torch, p, w_l, h_l, p_iter, pweight, None), want)


def test_static_scratch_entry_refuses_levels_beyond_its_array(gpu):
    torch = pytest.importorskip("torch")
    p = to_dis(gpu, base_orc(1, 1, psz=8, patove=.4))
    w_l, h_l = 3840, 2160      # 3 * 131 k cells + 518 k patches ints > DISB200_STATIC_SCRATCH_BYTES / 4
    n = gpu.patch_count(p, w_l, h_l)
    assert 3 * ((w_l + 16) // 8 + 2) * ((h_l + 16) // 8 + 2) + n > gpu.STATIC_SCRATCH_BYTES // 4
    d = torch.zeros(16, dtype=torch.float32, device="cuda")   # never dereferenced: the size check comes first
    with pytest.raises(gpu.DisError, match="scratch"):
        gpu.aggregate_dense_static(p, w_l, h_l, d.data_ptr(), d.data_ptr(), [d.data_ptr()] * 3, d.data_ptr(), 0)
    # the same level without the merge needs no scratch and is not refused on size; a missing array is
    with pytest.raises(gpu.DisError, match="null"):
        gpu.aggregate_dense_static(p, w_l, h_l, d.data_ptr(), 0, None, d.data_ptr(), 0)
