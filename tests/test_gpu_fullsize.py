"""-m gpu parity at BASELINE.json's full sizes (the oracle encodes 12 MP in ~0.3 s, so the
full-size cases are still compared bit for bit), size-independent properties of the fit, the
reference's own three PNG tests restated at the seam, and concurrent callers."""
import io
import threading

import numpy as np
import pytest

from image_compression_b200 import synth, CompressionParams, LearnedParams, FMT_RGBA8, FMT_RGB8

pytestmark = pytest.mark.gpu


def _psnr(a, b):
    d = a.astype(np.float64) - b.astype(np.float64)
    return 10 * np.log10(255.0 ** 2 / max(np.mean(d * d), 1e-12))


# ---- BASELINE configs[0]: 12 MP, -t 1048576 -q 0.25 -------------------------------------------
@pytest.mark.parametrize("seed", [0, 1])
def test_cfg1_12mp_fit_bit_exact(engine, oracle, seed):
    img = synth.natural_image(seed, 4000, 3000)
    assert np.array_equal(engine.jpg_coefficients(img), oracle.forward_coefficients(img))
    rc, ref, rq, rs, _, _ = oracle.jpg_fit(img, 0.25, 1048576)
    r = engine.jpg_fit(img, 0.25, 1048576)
    assert (r.status, r.best_q, r.scale) == (rc, rq, rs)
    assert r.data == ref
    assert len(r.data) <= 1048576
    # the stream is a real JPEG of that image (independent decoder)
    from PIL import Image
    dec = np.asarray(Image.open(io.BytesIO(r.data)).convert("RGB"))
    assert dec.shape == img.shape
    assert _psnr(dec, img) > 24.0


# ---- BASELINE configs[1]: 24 MP, 512 KB target ------------------------------------------------
def test_cfg2_24mp_fit_bit_exact(engine, oracle):
    img = synth.natural_image(100, 6000, 4000)
    rc, ref, rq, rs, _, _ = oracle.jpg_fit(img, 0.25, 524288)
    r = engine.jpg_fit(img, 0.25, 524288)
    assert (r.status, r.best_q, r.scale) == (rc, rq, rs)
    assert r.data == ref


# ---- BASELINE configs[4]: one 100 MP source (decoded 5774x4330 after s=2), 256 KB, -q 1.0 ------
def test_cfg5_100mp_full_sweep_bit_exact(engine, oracle):
    img = synth.natural_image(200, 5774, 4330)
    rc, ref, rq, rs, _, _ = oracle.jpg_fit(img, 1.0, 262144)
    r = engine.jpg_fit(img, 1.0, 262144)
    assert (r.status, r.best_q, r.scale) == (rc, rq, rs)
    if rc == 0:
        assert r.data == ref and len(r.data) <= 262144


def test_noise_24mp_stuffing_and_sizes(engine, oracle):
    """High-entropy input at the top of the quality range: the most 0xFF bytes per stream."""
    img = synth.noise_image(9, 2048, 1536)
    qs = [0.9921875, 0.75, 0.5, 0.2421875, 0.015625]
    got = engine.jpg_sizes(img, qs)
    assert list(got) == [oracle.encoded_size(img, q) for q in qs]
    data = engine.compress_jpg_to_stream(img, 0.9921875)
    assert data == oracle.encode(img, 0.9921875)
    assert data.count(b"\xff\x00") > 1000


# ---- properties that do not need the oracle -----------------------------------------------------
def test_fit_properties(engine):
    img = synth.natural_image(7, 1600, 1200)
    sizes = engine.jpg_sizes(img, [k / 128.0 for k in range(2, 128, 5)])
    assert all(int(b) >= int(a) for a, b in zip(sizes, sizes[1:]))        # monotone on this image
    target = int(sizes[6])
    a = engine.jpg_fit(img, 1.0, target)
    b = engine.jpg_fit(img, 1.0, target)
    assert a.status == 0 and a.data == b.data and a.best_q == b.best_q    # deterministic
    assert len(a.data) <= target
    # the chosen quality is a node of the dyadic tree and its explicit encode is the same stream
    assert a.best_q * 128 == int(a.best_q * 128)
    assert engine.compress_jpg_to_stream(img, a.best_q) == a.data
    # learned parameters reproduce it in one encode (core/ImageCompressionJpg.java:216-238)
    c = engine.jpg_fit(img, 1.0, target, cached=LearnedParams(a.best_q, a.scale))
    assert c.used_cache and c.data == a.data
    # a target below the header cannot be met at any scale (:120-121)
    assert engine.jpg_fit(img, 0.25, 600).status == 1


def test_seam_mirror_updates_the_cache(engine, tmp_path):
    img = synth.natural_image(8, 800, 600)
    params = CompressionParams(0.25, 0, 0, 0, 40000)
    cache = {}
    out = tmp_path / "a.jpg"
    assert engine.compress_jpg_with_target_size(img, 123456, out, params, cache) is True
    assert out.stat().st_size <= 40000 and len(cache) == 1
    first = out.read_bytes()
    (lp,) = cache.values()
    assert engine.compress_jpg_with_target_size(img, 123456, out, params, cache) is True
    assert engine.last_fit.used_cache and out.read_bytes() == first and list(cache.values()) == [lp]
    with pytest.raises(TypeError):
        engine.compress_jpg_with_target_size(None, 1, out, params, cache)


# ---- src/test/.../ImageCompressionPngTest.java:38-88, at the seam --------------------------------
def test_png_image_smaller_than_target_should_save_directly(engine):
    img = np.zeros((100, 100, 3), np.uint8)
    img[..., 0] = 255
    ok, out = engine.compress_png_with_target_size(img, "out.png", CompressionParams(0, 0, 100, 100, 0))
    assert ok is False and out is None


def test_png_image_larger_than_target_should_resize(engine, oracle):
    img = np.zeros((600, 800, 3), np.uint8)
    img[..., 2] = 255
    ok, out = engine.compress_png_with_target_size(img, "out.png", CompressionParams(0, 0, 100, 100, 0))
    assert ok is True
    assert out.shape == (75, 100, 3) and (out.shape[1] <= 100 or out.shape[0] <= 100)
    assert np.array_equal(out, oracle.resize(img, 100, 75))
    assert (out[..., 2] == 255).all() and (out[..., :2] == 0).all()


def test_png_null_inputs_should_throw(engine):
    img = np.zeros((8, 8, 3), np.uint8)
    p = CompressionParams(0, 0, 100, 100, 0)
    for args in [(None, "o", p), (img, None, p), (img, "o", None)]:
        with pytest.raises(TypeError):
            engine.compress_png_with_target_size(*args)


# ---- BASELINE configs[2]: 4K / 8K PNG pixels into the 1920 x 1920 box -----------------------------
@pytest.mark.parametrize("w,h,ch", [(3840, 2160, 4), (3840, 2160, 3), (7680, 4320, 4), (7680, 4320, 3)])
def test_cfg3_png_downscale(engine, oracle, w, h, ch):
    img = synth.natural_image(300 + ch, w, h, channels=ch)
    ok, out = engine.compress_png_with_target_size(img, "-", CompressionParams(0, 0, 1920, 1920, 0))
    assert ok and out.shape == (1080, 1920, ch)
    ofmt = oracle.FMT_RGBA8 if ch == 4 else oracle.FMT_RGB8
    assert np.array_equal(out, oracle.resize(img, 1920, 1080, fmt=ofmt))
    if ch == 3:          # stated tolerance: +-1 LSB against the exact (float64) bilinear sample
        f64 = oracle.resize(img, 1920, 1080, fmt=ofmt, f64=True)
        assert np.abs(out.astype(int) - f64.astype(int)).max() <= 1


# ---- input staging: the 100 MP source of configs[4] decimated like the reference's decoder does ----
def test_subsample_matches_the_decode_rule(engine, oracle):
    import image_compression_b200 as pkg
    assert pkg.subsample_factor(11548, 8660) == 2
    rng = np.random.default_rng(11)
    for (w, h, c, s) in [(1001, 777, 3, 2), (640, 480, 4, 4), (333, 251, 3, 3), (4097, 64, 3, 1)]:
        img = rng.integers(0, 256, (h, w, c), dtype=np.uint8)
        got = engine.subsample(img, s)
        assert np.array_equal(got, oracle.subsample(img, s)) and np.array_equal(got, img[::s, ::s])
    big = synth.natural_image(500, 5000, 3000)
    assert np.array_equal(engine.subsample(big, 2), big[::2, ::2])


# ---- CompressionBatch.java:68,77 -- the seams are called from many pool threads at once ----------
def test_concurrent_callers(engine, oracle):
    imgs = [synth.natural_image(400 + i, 640 + 16 * i, 480 + 8 * i) for i in range(8)]
    want = [oracle.jpg_fit(im, 0.5, 30000) for im in imgs]
    got = [None] * len(imgs)

    def work(i):
        for _ in range(3):
            got[i] = engine.jpg_fit(imgs[i], 0.5, 30000)

    ts = [threading.Thread(target=work, args=(i,)) for i in range(len(imgs))]
    for t in ts:
        t.start()
    for t in ts:
        t.join()
    for r, (rc, ref, rq, rs, _, _) in zip(got, want):
        assert (r.status, r.best_q, r.scale) == (rc, rq, rs)
        if rc == 0:
            assert r.data == ref
