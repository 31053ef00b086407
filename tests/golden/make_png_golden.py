"""Generates tests/golden/png/*.png and png_expected.npz: what cv2.imread (OpenCV 4.13, this container)
returns for each file with IMREAD_GRAYSCALE and IMREAD_COLOR.  disb200_read_image must reproduce these
arrays bit for bit (tests/test_io_cpu.py).  Run here only: cv2 does not travel to the GPU box.

Files written by cv2.imwrite: grey8, grey16, bgr8, bgra8, bgr16.  Files written by the small encoder below
(every filter type cycled over the rows, several IDAT chunks): palette 8/4/2/1 bit, grey 4/2/1 bit,
grey + alpha 8 bit.
"""
import os
import struct
import zlib

import cv2
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(HERE, "png")


def chunk(tag, body):
    return struct.pack(">I", len(body)) + tag + body + struct.pack(">I", zlib.crc32(tag + body) & 0xFFFFFFFF)


def paeth(a, b, c):
    p = a + b - c
    pa, pb, pc = abs(p - a), abs(p - b), abs(p - c)
    return a if (pa <= pb and pa <= pc) else (b if pb <= pc else c)


def encode(rows, w, h, depth, ctype, plte=None, bpp=1):
    """rows: list of bytes objects (packed scanlines).  Cycles filter types 0..4 over the rows."""
    raw = bytearray()
    prev = bytes(len(rows[0]))
    for y, row in enumerate(rows):
        ft = y % 5
        out = bytearray(len(row))
        for i, v in enumerate(row):
            a = row[i - bpp] if i >= bpp else 0
            b = prev[i]
            c = prev[i - bpp] if i >= bpp else 0
            pred = (0, a, b, (a + b) >> 1, paeth(a, b, c))[ft]
            out[i] = (v - pred) & 255
        raw.append(ft)
        raw += out
        prev = row
    z = zlib.compress(bytes(raw), 6)
    parts = [z[: len(z) // 3], z[len(z) // 3: 2 * len(z) // 3], z[2 * len(z) // 3:]]
    data = b"\x89PNG\r\n\x1a\n" + chunk(b"IHDR", struct.pack(">IIBBBBB", w, h, depth, ctype, 0, 0, 0))
    if plte is not None:
        data += chunk(b"PLTE", bytes(plte))
    for part in parts:
        data += chunk(b"IDAT", part)
    return data + chunk(b"IEND", b"")


def pack_bits(idx, depth):
    h, w = idx.shape
    rows = []
    for y in range(h):
        bits = 0
        nb = 0
        out = bytearray()
        for x in range(w):
            bits = (bits << depth) | int(idx[y, x])
            nb += depth
            if nb == 8:
                out.append(bits)
                bits = nb = 0
        if nb:
            out.append(bits << (8 - nb))
        rows.append(bytes(out))
    return rows


def main():
    os.makedirs(OUT, exist_ok=True)
    rng = np.random.default_rng(7)
    h, w = 23, 37
    yy, xx = np.mgrid[0:h, 0:w]
    smooth = ((np.sin(xx / 5.0) + np.cos(yy / 4.0)) * 60 + 128).clip(0, 255)
    files = {}
    g8 = (smooth + rng.integers(-20, 20, (h, w))).clip(0, 255).astype(np.uint8)
    files["grey8"] = g8
    files["grey16"] = (g8.astype(np.uint16) << 8) | rng.integers(0, 256, (h, w)).astype(np.uint16)
    bgr8 = np.stack([(smooth * k + rng.integers(0, 60, (h, w))).clip(0, 255) for k in (0.5, 0.8, 1.0)], -1).astype(np.uint8)
    files["bgr8"] = bgr8
    files["bgra8"] = np.concatenate([bgr8, rng.integers(0, 256, (h, w, 1)).astype(np.uint8)], -1)
    files["bgr16"] = (bgr8.astype(np.uint16) << 8) | rng.integers(0, 256, (h, w, 3)).astype(np.uint16)
    for name, arr in files.items():
        assert cv2.imwrite(os.path.join(OUT, name + ".png"), arr)
    for depth in (8, 4, 2, 1):
        ncol = 1 << depth
        plte = rng.integers(0, 256, ncol * 3).astype(np.uint8)
        idx = (g8.astype(int) * ncol // 256)
        rows = pack_bits(idx, depth) if depth < 8 else [bytes(idx[y].astype(np.uint8)) for y in range(h)]
        open(os.path.join(OUT, "pal%d.png" % depth), "wb").write(encode(rows, w, h, depth, 3, plte=plte))
        if depth < 8:
            open(os.path.join(OUT, "grey%d.png" % depth), "wb").write(encode(rows, w, h, depth, 0))
    ga = np.stack([g8, rng.integers(0, 256, (h, w)).astype(np.uint8)], -1)
    open(os.path.join(OUT, "greyalpha8.png"), "wb").write(encode([bytes(ga[y].tobytes()) for y in range(h)], w, h, 8, 4, bpp=2))
    exp = {}
    for f in sorted(os.listdir(OUT)):
        if f.endswith(".png"):
            p = os.path.join(OUT, f)
            exp[f[:-4] + "_grey"] = cv2.imread(p, cv2.IMREAD_GRAYSCALE)
            exp[f[:-4] + "_bgr"] = cv2.imread(p, cv2.IMREAD_COLOR)
            assert exp[f[:-4] + "_grey"] is not None, f
    np.savez_compressed(os.path.join(HERE, "png_expected.npz"), **exp)
    print("wrote", len(exp) // 2, "files; cv2", cv2.__version__)


if __name__ == "__main__":
    main()
