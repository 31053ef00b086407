"""Host-side I/O either side of the hot path: image decode standing in for
cv::imread (run_dense.cpp:209-210) and the asynchronous .flo / .pfm writers (run_dense.cpp:17-82).
No GPU needed: these entry points are plain host code inside libdisb200.so."""
import os

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PNG = os.path.join(ROOT, "tests", "golden", "png")
NAMES = sorted(f[:-4] for f in os.listdir(PNG) if f.endswith(".png"))


@pytest.fixture(scope="module")
def dis():
    import of_dis_b200 as d
    return d


@pytest.fixture(scope="module")
def expected():
    return np.load(os.path.join(ROOT, "tests", "golden", "png_expected.npz"))


@pytest.mark.parametrize("name", NAMES)
def test_png_decode_matches_cv2_imread(dis, expected, name):
    """Every colour type / bit depth / filter type, against what OpenCV 4.13 returned for the same file
    (tests/golden/make_png_golden.py): grey and BGR reads are bit-identical."""
    path = os.path.join(PNG, name + ".png")
    assert np.array_equal(dis.read_image(path, 1), expected[name + "_grey"])
    assert np.array_equal(dis.read_image(path, 3), expected[name + "_bgr"])


def test_read_image_dispatches_pnm_and_rejects_garbage(dis, tmp_path):
    img = np.arange(6 * 5, dtype=np.uint8).reshape(5, 6)
    p = tmp_path / "a.pgm"
    p.write_bytes(b"P5\n6 5\n255\n" + img.tobytes())
    assert np.array_equal(dis.read_image(str(p), 1), img)
    assert np.array_equal(dis.read_image(str(p), 3), np.repeat(img[..., None], 3, -1))
    bad = tmp_path / "bad.png"
    bad.write_bytes(b"\x89PNG\r\n\x1a\n" + b"\0" * 40)
    with pytest.raises(dis.DisError):
        dis.read_image(str(bad), 1)
    with pytest.raises(dis.DisError):
        dis.read_image(str(tmp_path / "missing.png"), 1)
    # truncated IDAT stream
    good = open(os.path.join(PNG, "bgr8.png"), "rb").read()
    cut = tmp_path / "cut.png"
    cut.write_bytes(good[: len(good) // 2])
    with pytest.raises(dis.DisError):
        dis.read_image(str(cut), 3)


def test_interlaced_png_is_reported_unsupported(dis, tmp_path):
    good = bytearray(open(os.path.join(PNG, "grey8.png"), "rb").read())
    import struct
    import zlib
    assert good[12:16] == b"IHDR"
    good[28] = 1                                             # interlace method
    good[29:33] = struct.pack(">I", zlib.crc32(bytes(good[12:29])) & 0xFFFFFFFF)
    p = tmp_path / "adam7.png"
    p.write_bytes(bytes(good))
    with pytest.raises(dis.DisError) as e:
        dis.read_image(str(p), 1)
    assert e.value.code == 4          # DISB200_EUNSUPPORTED


def test_readers_survive_hostile_headers(dis, tmp_path):
    """The readers sit behind a C ABI and are fed user-listed files: a header that declares an absurd size, a truncated
    stream or an endless digit run must come back as an error code, never as an allocation failure or a crash."""
    import struct
    import zlib
    good = bytearray(open(os.path.join(PNG, "grey8.png"), "rb").read())
    huge = bytearray(good)
    huge[16:24] = struct.pack(">II", 1 << 20, 1 << 20)      # IHDR width, height: 2^40 pixels
    huge[29:33] = struct.pack(">I", zlib.crc32(bytes(huge[12:29])) & 0xFFFFFFFF)
    (tmp_path / "huge.png").write_bytes(bytes(huge))
    i = bytes(good).index(b"IDAT")
    (tmp_path / "cut.png").write_bytes(bytes(good[:i + 4 + 10]) + bytes(good[-12:]))   # IDAT body cut short (chunk length now lies)
    (tmp_path / "digits.pgm").write_bytes(b"P5\n" + b"9" * 4000 + b" 4\n255\n" + bytes(16))
    (tmp_path / "huge.pgm").write_bytes(b"P5\n1000000 1000000\n255\n" + bytes(16))
    for name in ("huge.png", "cut.png", "digits.pgm", "huge.pgm"):
        with pytest.raises(dis.DisError) as e:
            dis.read_image(str(tmp_path / name), 1)
        assert e.value.code in (3, 4, 5), (name, e.value.code)     # EIO / EUNSUPPORTED / ENOMEM


def test_writer_pool_writes_the_same_bytes_as_the_synchronous_writers(dis, tmp_path):
    rng = np.random.default_rng(0)
    flows = [rng.standard_normal((17 + i, 31, 2)).astype(np.float32) for i in range(24)]
    disps = [rng.standard_normal((9, 14 + i)).astype(np.float32) for i in range(8)]
    with dis.Writer(threads=4, max_queued_bytes=3 * flows[0].nbytes) as w:      # small queue: submit must block, not drop
        for i, f in enumerate(flows):
            w.submit_flo(str(tmp_path / ("a%02d.flo" % i)), f)
            f[:] = 0                                                           # the pool owns a copy
        for i, d in enumerate(disps):
            w.submit_pfm(str(tmp_path / ("d%02d.pfm" % i)), d)
        w.wait()
        rng = np.random.default_rng(0)
        for i in range(24):
            f = rng.standard_normal((17 + i, 31, 2)).astype(np.float32)
            dis.save_flo(str(tmp_path / "ref.flo"), f)
            assert (tmp_path / ("a%02d.flo" % i)).read_bytes() == (tmp_path / "ref.flo").read_bytes()
        for i, d in enumerate(disps):
            dis.save_pfm(str(tmp_path / "ref.pfm"), d)
            assert (tmp_path / ("d%02d.pfm" % i)).read_bytes() == (tmp_path / "ref.pfm").read_bytes()
        # an unwritable path surfaces at wait(), once
        w.submit_flo(str(tmp_path / "no_such_dir" / "x.flo"), flows[0])
        with pytest.raises(dis.DisError):
            w.wait()
        w.wait()
