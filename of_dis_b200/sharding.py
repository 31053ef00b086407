"""Frame-pair sharding across the GPUs of one box.

Frame pairs are independent units: rank r of `world` owns a contiguous block of pairs and there
is no data-path collective.  For a video (C5) consecutive pairs share a frame, so a contiguous
block lets each interior frame's pyramid be built once per GPU.
"""
from __future__ import annotations


def shard_range(n_pairs: int, rank: int, world: int):
    """Contiguous, balanced [lo, hi) block of pair indices for `rank` (first `n % world` ranks get one more)."""
    if world < 1 or not (0 <= rank < world) or n_pairs < 0:
        raise ValueError("bad shard request")
    base, extra = divmod(n_pairs, world)
    lo = rank * base + min(rank, extra)
    hi = lo + base + (1 if rank < extra else 0)
    return lo, hi


def frames_for_pairs(lo: int, hi: int):
    """Video: pair i uses frames (i, i+1); the block [lo, hi) needs frames [lo, hi]."""
    return (lo, hi + 1) if hi > lo else (lo, lo)
