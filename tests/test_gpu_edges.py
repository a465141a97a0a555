"""-m gpu parity on the edge cases of the seams: padded row pitch, BGR / ABGR rasters, the
alternative [JDK-mem] profile, tiny images, capacity and argument errors, transparent pixels."""
import ctypes as C

import numpy as np
import pytest

import image_compression_b200 as pkg
from image_compression_b200 import api, synth, FMT_RGB8, FMT_BGR8, FMT_RGBA8, FMT_ABGR8

pytestmark = pytest.mark.gpu


def _raw_fit(engine, buf, w, h, pitch, fmt, q, target):
    out = np.empty(target + 16, np.uint8)
    n, bq, sc, uc = C.c_size_t(0), C.c_float(0), C.c_double(0), C.c_int(0)
    rc = engine._L.icb_jpg_fit(engine._h, buf.ctypes.data, w, h, pitch, fmt, C.c_float(q), target, 0, C.c_float(0),
                               C.c_double(1.0), out.ctypes.data, out.size, C.byref(n), C.byref(bq), C.byref(sc),
                               C.byref(uc))
    return rc, out[: n.value].tobytes(), bq.value, sc.value


@pytest.mark.parametrize("w,h,pad", [(333, 251, 5), (640, 480, 64), (17, 17, 1)])
def test_padded_pitch_rows(engine, oracle, w, h, pad):
    """BufferedImage rasters may carry a scanline stride larger than w*3."""
    img = synth.natural_image(40 + w, w, h)
    pitch = w * 3 + pad
    buf = np.full((h, pitch), 0xA5, np.uint8)
    buf[:, : w * 3] = img.reshape(h, w * 3)
    target = oracle.encoded_size(img, 0.2) + 7
    rc, ref, rq, rs, _, _ = oracle.jpg_fit(img, 0.5, target)
    got = _raw_fit(engine, buf, w, h, pitch, FMT_RGB8, 0.5, target)
    assert got == (rc, ref, rq, rs)


def test_bgr_raster_full_fit(engine, oracle):
    """TYPE_3BYTE_BGR is what the ImageIO JPEG reader hands the seam."""
    img = synth.natural_image(61, 801, 603)
    bgr = np.ascontiguousarray(img[..., ::-1])
    target = 60000
    rc, ref, rq, rs, _, _ = oracle.jpg_fit(img, 0.75, target)
    r = engine.jpg_fit(bgr, 0.75, target, fmt=FMT_BGR8)
    assert (r.status, r.best_q, r.scale) == (rc, rq, rs) and r.data == ref
    assert engine.compress_jpg_to_stream(bgr, 0.33, fmt=FMT_BGR8) == oracle.encode(img, 0.33)


def test_alternative_profile_ijg_tables_jfif_101(oracle):
    """The [JDK-mem] facts live behind one switch: IJG integer scaling + JFIF 1.01."""
    img = synth.natural_image(62, 320, 200)
    with pkg.Engine(slots_per_gpu=1) as eng:
        eng.set_profile(1, 1)
        for q in (0.1, 0.37, 0.5, 0.93):
            assert eng.compress_jpg_to_stream(img, q) == oracle.encode(img, q, prof=oracle.IJG)
        rc, ref, rq, rs, _, _ = oracle.jpg_fit(img, 1.0, 9000, prof=oracle.IJG)
        r = eng.jpg_fit(img, 1.0, 9000)
        assert (r.status, r.best_q, r.scale) == (rc, rq, rs) and r.data == ref
        eng.set_profile(0, 2)
        assert eng.compress_jpg_to_stream(img, 0.37) == oracle.encode(img, 0.37)


@pytest.mark.parametrize("w,h", [(1, 1), (2, 3), (7, 7), (8, 1), (1, 40), (16, 9)])
def test_tiny_images_fit_and_no_fit(engine, oracle, w, h):
    img = synth.noise_image(70 + w * h, w, h)
    for target in (700, 640, 100):
        rc, ref, rq, rs, _, _ = oracle.jpg_fit(img, 0.25, target)
        r = engine.jpg_fit(img, 0.25, target)
        assert r.status == rc
        if rc == 0:
            assert (r.best_q, r.scale) == (rq, rs) and r.data == ref


@pytest.mark.parametrize("fmt,ofmt_name", [(FMT_ABGR8, "FMT_ABGR8"), (FMT_RGBA8, "FMT_RGBA8")])
def test_resize_with_alpha_edges(engine, oracle, fmt, ofmt_name):
    """IntArgbPre round trip: fully transparent, fully opaque and fractional alpha."""
    rng = np.random.default_rng(80)
    img = rng.integers(0, 256, (190, 263, 4), dtype=np.uint8)
    a = 0 if fmt == FMT_ABGR8 else 3
    img[:40, :, a] = 0
    img[40:90, :, a] = 255
    img[90:120, :, a] = 1
    for dw, dh in [(131, 95), (100, 75), (263, 190), (1, 1)]:
        got = engine.resize_to(img, dw, dh, fmt)
        assert np.array_equal(got, oracle.resize(img, dw, dh, fmt=getattr(oracle, ofmt_name)))


def test_capacity_and_argument_errors(engine):
    img = synth.natural_image(90, 64, 48)
    L, h = engine._L, engine._h
    n = C.c_size_t(0)
    small = np.empty(100, np.uint8)
    rc = L.icb_jpg_encode(h, img.ctypes.data, 64, 48, 64 * 3, FMT_RGB8, C.c_float(0.9), small.ctypes.data, 100,
                          C.byref(n))
    assert rc == api.ICB_ERR_CAPACITY and n.value > 100             # *out_len = bytes needed
    big = np.empty(n.value, np.uint8)
    assert L.icb_jpg_encode(h, img.ctypes.data, 64, 48, 64 * 3, FMT_RGB8, C.c_float(0.9), big.ctypes.data, big.size,
                            C.byref(n)) == api.ICB_OK
    # alpha / gray rasters do not go through the JPG path (the caller keeps the stock writer)
    assert L.icb_jpg_encode(h, img.ctypes.data, 64, 48, 64 * 4, FMT_RGBA8, C.c_float(0.9), big.ctypes.data, big.size,
                            C.byref(n)) == api.ICB_ERR_UNSUPPORTED
    assert L.icb_jpg_encode(h, img.ctypes.data, 64, 48, 10, FMT_RGB8, C.c_float(0.9), big.ctypes.data, big.size,
                            C.byref(n)) == api.ICB_ERR_INVALID     # pitch smaller than a row
    assert L.icb_jpg_encode(h, img.ctypes.data, 0, 48, 192, FMT_RGB8, C.c_float(0.9), big.ctypes.data, big.size,
                            C.byref(n)) == api.ICB_ERR_INVALID
    assert L.icb_jpg_encode(h, None, 64, 48, 192, FMT_RGB8, C.c_float(0.9), big.ctypes.data, big.size,
                            C.byref(n)) == api.ICB_ERR_INVALID
    assert L.icb_wait(h, 0x7fff0001, None) == api.ICB_ERR_TICKET
    assert (engine._L.icb_last_error(h) or b"") != b""


def test_quality_range_like_the_jdk(engine, oracle):
    """ImageWriteParam.setCompressionQuality accepts [0, 1] (JPEG.convertToLinearQuality then clamps the
    low end) and throws IllegalArgumentException outside it: the seam accepts exactly what Java accepts."""
    img = synth.natural_image(91, 96, 64)
    for q in (0.0, 0.005, 1.0):
        assert engine.compress_jpg_to_stream(img, q) == oracle.encode(img, q)
    for q in (1.5, -0.1, float("nan")):
        with pytest.raises(api.IcbError) as e:
            engine.compress_jpg_to_stream(img, q)
        assert e.value.code == api.ICB_ERR_INVALID
    for kw in (dict(init_q=1.5), dict(init_q=float("nan")), dict(init_q=0.25, cached=pkg.LearnedParams(1.2, 1.0)),
               dict(init_q=0.25, cached=pkg.LearnedParams(float("nan"), 1.0))):
        with pytest.raises(api.IcbError) as e:
            engine.jpg_fit(img, kw["init_q"], 5000, kw.get("cached"))
        assert e.value.code == api.ICB_ERR_INVALID
    # a non-positive initial quality is legal Java: every midpoint is below 0.01, nothing is encoded, no scale fits
    assert engine.jpg_fit(img, 0.0, 5000).status == api.ICB_NO_FIT
    assert engine.jpg_fit(img, -1.0, 5000).status == api.ICB_NO_FIT
