"""Seeded synthetic warped-texture frame pairs (SURVEY.md 8(d) "Synthetic inputs").

The reference ships no sample data, so every test and the bench
run on pairs made here: a multi-octave band-limited noise texture on a canvas
100 px larger than the frame on each side; frame A is the centre crop, frame B
is the canvas backward-warped (bicubic) by a known smooth field and quantised to
u8.  Ground truth is the field itself (approximate at the sub-0.1 px level
because B is resampled).  Host-side numpy/scipy only: this is input generation,
not part of the measured path.
"""
from __future__ import annotations

import numpy as np
from scipy import ndimage

MARGIN = 100


def _texture(h: int, w: int, rng: np.random.Generator) -> np.ndarray:
    """Multi-octave noise: octave cell sizes 1,2,4,8,16,32 px, amplitudes 20..45."""
    acc = np.zeros((h, w), np.float64)
    cells = [1, 2, 4, 8, 16, 32]
    amps = np.linspace(20.0, 45.0, len(cells))
    for cell, amp in zip(cells, amps):
        gh, gw = -(-h // cell) + 3, -(-w // cell) + 3
        g = rng.standard_normal((gh, gw))
        if cell == 1:
            up = g[:h, :w]
        else:
            up = ndimage.zoom(g, cell, order=3, mode="nearest")[:h, :w]
        up = up / (up.std() + 1e-12)
        acc += amp * up / len(cells) * 2.0
    return acc + 128.0


def of_field(h: int, w: int):
    """Smooth optical-flow field, roughly +-8 px."""
    y, x = np.mgrid[0:h, 0:w].astype(np.float64)
    cx, cy = w / 2.0, h / 2.0
    u = 3 + 0.01 * (x - cx) - 0.008 * (y - cy) + 2 * np.sin(y / 40.0)
    v = -1.5 + 0.006 * (x - cx) + 0.01 * (y - cy) + 1.5 * np.cos(x / 55.0)
    return u, v


def de_field(h: int, w: int):
    """Stereo disparity 4..10 px, negative in the left view (patch.cpp:187-188)."""
    y, x = np.mgrid[0:h, 0:w].astype(np.float64)
    u = -(4 + 6 * (0.5 + 0.5 * np.sin(x / 90.0) * np.cos(y / 70.0)))
    return u, np.zeros_like(u)


def make_pair(h: int, w: int, channels: int = 1, mode: str = "OF", seed: int = 0, t: float = 1.0):
    """Return (A, B, gt) with A,B uint8 arrays (h,w) or (h,w,3) and gt float32 (h,w,nop)."""
    rng = np.random.default_rng(seed)
    H, W = h + 2 * MARGIN, w + 2 * MARGIN
    u, v = of_field(h, w) if mode == "OF" else de_field(h, w)
    u, v = u * t, v * t
    yy, xx = np.mgrid[0:h, 0:w].astype(np.float64)
    a_ch, b_ch = [], []
    for _ in range(channels):
        canvas = _texture(H, W, rng)
        a = canvas[MARGIN:MARGIN + h, MARGIN:MARGIN + w]
        # A(x) = B(x + flow(x)): build B by a fixed-point inverse of the smooth field.
        # B(p) = canvas(p - flow(q)), with q + flow(q) = p.
        qx, qy = xx.copy(), yy.copy()
        for _ in range(6):
            fu = ndimage.map_coordinates(u, [qy, qx], order=1, mode="nearest")
            fv = ndimage.map_coordinates(v, [qy, qx], order=1, mode="nearest")
            qx, qy = xx - fu, yy - fv
        b = ndimage.map_coordinates(canvas, [qy + MARGIN, qx + MARGIN], order=3, mode="nearest")
        a_ch.append(np.clip(np.rint(a), 0, 255).astype(np.uint8))
        b_ch.append(np.clip(np.rint(b), 0, 255).astype(np.uint8))
    if channels == 1:
        A, B = a_ch[0], b_ch[0]
    else:
        A, B = np.stack(a_ch, -1), np.stack(b_ch, -1)
    gt = np.stack([u, v], -1).astype(np.float32) if mode == "OF" else u[..., None].astype(np.float32)
    return np.ascontiguousarray(A), np.ascontiguousarray(B), gt


def make_sequence(h: int, w: int, n_frames: int, seed: int = 0):
    """Video for C5: frame t = canvas warped by t*field/4; returns list of u8 frames."""
    rng = np.random.default_rng(seed)
    H, W = h + 2 * MARGIN, w + 2 * MARGIN
    canvas = _texture(H, W, rng)
    u, v = of_field(h, w)
    yy, xx = np.mgrid[0:h, 0:w].astype(np.float64)
    frames = []
    for t in range(n_frames):
        s = t / 4.0
        f = ndimage.map_coordinates(canvas, [yy - s * v + MARGIN, xx - s * u + MARGIN], order=3, mode="nearest")
        frames.append(np.clip(np.rint(f), 0, 255).astype(np.uint8))
    return frames
