"""-m gpu parity tests proper: the CUDA path through the C ABI against the CPU oracle,
bit-exact for coefficients, stream sizes, JFIF bytes and chosen quality."""
import numpy as np
import pytest

from image_compression_b200 import synth, LearnedParams, CompressionParams

pytestmark = pytest.mark.gpu

SIZES = [(1, 1), (8, 8), (16, 16), (17, 17), (40, 56), (15, 33), (33, 15), (255, 31), (257, 49),
         (333, 251), (640, 480), (1023, 767), (4000, 24), (24, 1000)]


def _img(kind, w, h, idx=0):
    if kind == "nat":
        return synth.natural_image(idx + w * 7 + h, w, h)
    if kind == "noise":
        return synth.noise_image(idx + w + h, w, h)
    return synth.pattern_image(kind, w, h)


@pytest.mark.parametrize("w,h", SIZES)
@pytest.mark.parametrize("kind", ["nat", "noise", "checker"])
def test_transform_coefficients_exact(engine, oracle, w, h, kind):
    img = _img(kind, w, h)
    got = engine.jpg_coefficients(img)
    ref = oracle.forward_coefficients(img)
    assert got.shape == ref.shape
    assert np.array_equal(got, ref)


def test_transform_bgr(engine, oracle):
    img = synth.natural_image(5, 301, 203)
    bgr = np.ascontiguousarray(img[..., ::-1])
    from image_compression_b200 import FMT_BGR8
    assert np.array_equal(engine.jpg_coefficients(bgr, fmt=FMT_BGR8), oracle.forward_coefficients(img))


@pytest.mark.parametrize("w,h", SIZES)
@pytest.mark.parametrize("q", [0.015625, 0.1, 0.2421875, 0.5, 0.9, 1.0])
def test_encode_bytes_exact(engine, oracle, w, h, q):
    img = _img("nat", w, h)
    assert engine.compress_jpg_to_stream(img, q) == oracle.encode(img, q)


@pytest.mark.parametrize("kind,q", [("noise", 1.0), ("noise", 0.97), ("noise", 0.6), ("checker", 0.95),
                                    ("solid", 0.5), ("gradient", 0.8)])
def test_encode_stuffing_stress(engine, oracle, kind, q):
    img = _img(kind, 515, 389)
    ref = oracle.encode(img, q)
    got = engine.compress_jpg_to_stream(img, q)
    assert len(got) == len(ref)
    assert got == ref


def test_sizes_all_candidates(engine, oracle):
    img = synth.natural_image(11, 801, 603)
    qs = [k / 128.0 for k in range(2, 128, 3)]
    got = engine.jpg_sizes(img, qs)
    ref = [oracle.encoded_size(img, q) for q in qs]
    assert list(got) == ref


@pytest.mark.parametrize("init_q", [0.25, 0.5, 0.75, 1.0])
@pytest.mark.parametrize("frac", [0.02, 0.3, 0.6, 0.9, 1.2, 5.0])
def test_fit_search_matches_reference_loop(engine, oracle, init_q, frac):
    img = synth.natural_image(21, 640, 480)
    full = oracle.encoded_size(img, init_q)
    target = int(full * frac)
    rc, ref_bytes, ref_q, ref_scale, _, _ = oracle.jpg_fit(img, init_q, target)
    r = engine.jpg_fit(img, init_q, target)
    assert r.status == rc
    if rc == 0:
        assert r.best_q == ref_q and r.scale == ref_scale
        assert r.data == ref_bytes


def test_fit_cached_params(engine, oracle):
    img = synth.natural_image(22, 512, 384)
    target = oracle.encoded_size(img, 0.2) + 10
    for cached in [(0.2, 1.0), (0.2, 0.85), (0.9, 1.0), (0.9, 0.5)]:
        rc, ref_bytes, ref_q, ref_scale, ref_used, _ = oracle.jpg_fit(img, 0.25, target, cached=cached)
        r = engine.jpg_fit(img, 0.25, target, cached=LearnedParams(*cached))
        assert (r.status, r.used_cache, r.best_q, r.scale) == (rc, bool(ref_used), ref_q, ref_scale)
        assert r.data == ref_bytes


@pytest.mark.parametrize("fmt_c", [3, 4])
@pytest.mark.parametrize("src,dst", [((800, 600), (100, 75)), ((640, 480), (320, 240)), ((1001, 777), (850, 660)),
                                     ((333, 251), (83, 62)), ((64, 64), (1, 1))])
def test_resize_exact(engine, oracle, src, dst, fmt_c):
    img = synth.natural_image(31, src[0], src[1], channels=fmt_c)
    from image_compression_b200 import FMT_RGB8, FMT_RGBA8
    fmt = FMT_RGB8 if fmt_c == 3 else FMT_RGBA8
    ofmt = oracle.FMT_RGB8 if fmt_c == 3 else oracle.FMT_RGBA8
    got = engine.resize_to(img, dst[0], dst[1], fmt)
    ref = oracle.resize(img, dst[0], dst[1], fmt=ofmt)
    assert np.array_equal(got, ref)
    if fmt_c == 3:      # the stated tolerance: +-1 LSB against the exact (float64) bilinear sample
        f64 = oracle.resize(img, dst[0], dst[1], fmt=ofmt, f64=True)
        assert np.abs(got.astype(int) - f64.astype(int)).max() <= 1
