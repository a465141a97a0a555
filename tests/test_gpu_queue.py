"""-m gpu tests of the queue side of the boundary (SURVEY 8b submit/wait/poll, 8e work queues):
pageable and page-locked callers, non-blocking poll, queued resize / PNG jobs, the extra raster
formats of tools/ImageTools.java:12-15, the warm-cache mix of BASELINE configs[3], and run_shard
with several jobs in flight pulled from one cursor.  Every result is compared with the oracle."""
import threading
import time

import numpy as np
import pytest

import image_compression_b200 as pkg
from image_compression_b200 import api, batch, synth
from image_compression_b200 import FMT_RGB8, FMT_RGBA8, FMT_ABGR8, FMT_GRAY8, FMT_ARGB32, FMT_XRGB32, LearnedParams

pytestmark = pytest.mark.gpu


def test_pageable_and_pinned_callers_agree(engine, oracle):
    """A numpy array is pageable memory (like a Java heap array copied by FFM): the library stages it in
    slices; a buffer from icb_alloc_pinned is copied as it is.  Same bytes either way, odd pitch included."""
    img = synth.natural_image(810, 1237, 911)
    rc, ref, rq, rs, _, _ = oracle.jpg_fit(img, 0.5, 70000)
    pin = engine.alloc_pinned(img.shape)
    pin[...] = img
    out_pin = engine.alloc_pinned((70016,))
    out_pg = np.zeros(70016, np.uint8)
    for src in (img, pin):
        for out in (out_pg, out_pin):
            out[...] = 0
            r = engine.wait(engine.submit_jpg_fit(src, 0.5, 70000, out))
            assert (r.status, r.best_q, r.scale) == (rc, rq, rs)
            assert out[: r.out_len].tobytes() == ref


def test_poll_never_blocks_and_finishes_multi_phase_jobs(engine, oracle):
    """icb_poll moves a fit through its phases (cached miss -> search -> next scale) without icb_wait."""
    img = synth.natural_image(811, 801, 603)
    target = 2500                                        # too small at scale 1: the 0.85 loop runs
    rc, ref, rq, rs, _, _ = oracle.jpg_fit(img, 0.25, target, cached=(0.9, 1.0))
    out = np.zeros(target + 16, np.uint8)
    t = engine.submit_jpg_fit(img, 0.25, target, out, cached=LearnedParams(0.9, 1.0))
    t0 = time.perf_counter()
    polls = 0
    while not engine.poll(t):
        polls += 1
        assert time.perf_counter() - t0 < 30
    assert engine.poll(t)                                # stays done
    r = engine.wait(t)
    assert (r.status, r.best_q, r.scale, r.used_cache) == (rc, rq, rs, False)
    if rc == 0:
        assert out[: r.out_len].tobytes() == ref
    with pytest.raises(api.IcbError):                    # the ticket is spent
        engine.poll(t)


@pytest.mark.parametrize("fmt,ofmt,c", [(FMT_GRAY8, "FMT_GRAY8", 1), (FMT_ARGB32, "FMT_ARGB32", 4),
                                        (FMT_XRGB32, "FMT_XRGB32", 4), (FMT_RGB8, "FMT_RGB8", 3),
                                        (FMT_RGBA8, "FMT_RGBA8", 4), (FMT_ABGR8, "FMT_ABGR8", 4)])
def test_resize_every_raster_format_and_scale(engine, oracle, fmt, ofmt, c):
    """tools/ImageTools.java:12-15: byte rasters, TYPE_BYTE_GRAY and the packed-int fallbacks.  Scales from the
    bottom of the JPG scale loop (0.1) to mild upscaling, ragged sizes: every destination tile width of the kernel."""
    rng = np.random.default_rng(820 + fmt)
    for (w, h) in [(1013, 517), (64, 48), (2051, 33), (5, 3)]:
        img = rng.integers(0, 256, (h, w, c), dtype=np.uint8)
        if c == 4:
            a = 0 if fmt == FMT_ABGR8 else 3
            img[: h // 3, :, a] = 255
            img[h // 3: h // 2, : w // 2, a] = 0
        for scale in (0.1, 0.12, 0.25, 0.37, 0.5, 0.7225, 0.85, 0.999, 1.3):
            dw, dh = pkg.resize_dims(w, h, scale)
            got = engine.resize_to(img, dw, dh, fmt)
            ref = oracle.resize(img, dw, dh, fmt=getattr(oracle, ofmt))
            assert np.array_equal(got, ref), (w, h, scale, fmt)


def _premul_bilinear_f64(img, dw, dh, a_idx):
    """float64 pixel-centre bilinear on premultiplied colour (what Java2D's IntArgbPre loop approximates)."""
    h, w, c = img.shape
    f = img.astype(np.float64)
    if a_idx is not None:
        al = f[..., a_idx:a_idx + 1] / 255.0
        pm = f * al
        pm[..., a_idx] = f[..., a_idx]
    else:
        pm = f
    sy = (np.arange(dh) + 0.5) * h / dh - 0.5
    sx = (np.arange(dw) + 0.5) * w / dw - 0.5
    y0 = np.floor(sy); fy = (sy - y0)[:, None, None]
    x0 = np.floor(sx); fx = (sx - x0)[None, :, None]
    y0i = np.clip(y0.astype(int), 0, h - 1); y1i = np.clip(y0.astype(int) + 1, 0, h - 1)
    x0i = np.clip(x0.astype(int), 0, w - 1); x1i = np.clip(x0.astype(int) + 1, 0, w - 1)
    top = pm[y0i][:, x0i] * (1 - fx) + pm[y0i][:, x1i] * fx
    bot = pm[y1i][:, x0i] * (1 - fx) + pm[y1i][:, x1i] * fx
    return top * (1 - fy) + bot * fy


@pytest.mark.parametrize("shape", [((800, 600), (100, 75)), ((640, 480), (320, 240)), ((1001, 777), (850, 660)),
                                   ((3840, 2160), (1920, 1080))])
def test_resize_alpha_tolerance_against_float64_premultiplied(engine, shape):
    """north_star: PNG pixels within +-1 LSB of the reference's Java2D result -- that is what the bit-exact
    comparison with the oracle's restatement of the Java2D loop checks.  THIS test measures how far that loop
    itself sits from the exact (float64) bilinear sample: its 8-bit tap fractions move a sample by up to
    255 * 2 / 256 = 2 LSB where neighbouring pixels differ by the full range.  Measured here (r02): <= 1.0 on
    the RGB channels of natural-statistics images, 1.50 on the synthetic alpha texture of an 0.85 scale.  With
    alpha the loop interpolates PREMULTIPLIED colour (IntArgbPre) and divides back, so colour is compared
    where the interpolation happens (colour x alpha / 255, plus the half LSB of each tap's mul8 rounding)."""
    (w, h), (dw, dh) = shape
    img = synth.natural_image(830, w, h, channels=4)
    got = engine.resize_to(img, dw, dh, FMT_RGBA8).astype(np.float64)
    ref = _premul_bilinear_f64(img, dw, dh, 3)
    assert np.abs(got[..., 3] - ref[..., 3]).max() <= 2.0
    got_pm = got[..., :3] * got[..., 3:4] / 255.0
    assert np.abs(got_pm - ref[..., :3]).max() <= 2.5
    opaque = ref[..., 3] >= 254.999
    if opaque.any():
        assert np.abs(got[..., :3][opaque] - ref[..., :3][opaque]).max() <= 1.5
    rgb = engine.resize_to(np.ascontiguousarray(img[..., :3]), dw, dh, FMT_RGB8).astype(np.float64)
    assert np.abs(rgb - _premul_bilinear_f64(img[..., :3], dw, dh, None)).max() <= 1.5


def test_queued_resize_and_png_jobs(engine, oracle):
    """icb_submit_resize / icb_submit_png_fit: the PNG seam in the queue, several jobs in flight."""
    imgs = [synth.natural_image(840 + i, 2100 + 64 * i, 1400 + 32 * i, channels=4) for i in range(3)]
    jobs = []
    for im in imgs:
        dst = np.zeros(1920 * 1920 * 4, np.uint8)
        t = engine.submit_png_fit(im, 1920, 1920, dst)
        assert t is not None
        jobs.append((t, dst, im))
    for (t, dw, dh), dst, im in jobs:
        r = engine.wait(t)
        assert r.status == 0
        assert (dw, dh) == pkg.png_target_dims(im.shape[1], im.shape[0], 1920, 1920)[:2]
        assert np.array_equal(dst[: dw * dh * 4].reshape(dh, dw, 4), oracle.resize(im, dw, dh, fmt=oracle.FMT_RGBA8))
    small = synth.natural_image(845, 640, 480, channels=4)
    assert engine.submit_png_fit(small, 1920, 1920, np.zeros(16, np.uint8)) is None      # :49-53 returns false


def test_cfg4_mix_cached_hits_misses_and_png_interleaved(engine, oracle):
    """BASELINE configs[3]: a list that mixes JPG pixels whose learned parameters hit, JPG pixels whose
    learned parameters no longer fit (fall back to the search, core/ImageCompressionJpg.java:82-89) and PNG
    pixels, all in flight together; every result against the oracle."""
    target = 40000
    jpgs = [synth.natural_image(850 + i, 900 + 40 * i, 700 + 24 * i) for i in range(4)]
    pngs = [synth.natural_image(860 + i, 2400, 1500 + 100 * i, channels=4) for i in range(2)]
    learned = []
    for im in jpgs:                                      # warm the cache: one search per key
        r = engine.jpg_fit(im, 0.5, target)
        assert r.status == 0
        learned.append(LearnedParams(r.best_q, r.scale))
    plan = []
    for rep in range(3):
        for i, im in enumerate(jpgs):
            hit = (i + rep) % 2 == 0
            lp = learned[i] if hit else LearnedParams(min(1.0, learned[i].quality + 0.4), 1.0)   # stale: overshoots
            plan.append(("jpg", im, lp))
            if i < len(pngs):
                plan.append(("png", pngs[i], None))
    inflight = []
    results = []

    def collect():
        kind, im, lp, t, out = inflight.pop(0)
        results.append((kind, im, lp, engine.wait(t if kind == "jpg" else t[0]), out, t))

    for kind, im, lp in plan:
        if len(inflight) == 3:
            collect()
        if kind == "jpg":
            out = np.zeros(target + 16, np.uint8)
            inflight.append((kind, im, lp, engine.submit_jpg_fit(im, 0.5, target, out, cached=lp), out))
        else:
            out = np.zeros(1920 * 1920 * 4, np.uint8)
            inflight.append((kind, im, lp, engine.submit_png_fit(im, 1920, 1920, out), out))
    while inflight:
        collect()
    hits = misses = 0
    for kind, im, lp, r, out, t in results:
        if kind == "jpg":
            rc, ref, rq, rs, uc, _ = oracle.jpg_fit(im, 0.5, target, cached=(lp.quality, lp.scale))
            assert (r.status, r.best_q, r.scale, int(r.used_cache)) == (rc, rq, rs, int(uc))
            assert out[: r.out_len].tobytes() == ref
            hits += int(r.used_cache)
            misses += int(not r.used_cache)
        else:
            _, dw, dh = t
            assert np.array_equal(out[: dw * dh * 4].reshape(dh, dw, 4), oracle.resize(im, dw, dh, fmt=oracle.FMT_RGBA8))
    assert hits > 0 and misses > 0


def test_run_shard_keeps_jobs_in_flight_from_one_cursor(engine, oracle):
    """batch.run_shard: several caller threads pull lines from ONE shared cursor and each keeps
    jobs in flight; counters and outputs equal the oracle's, whatever the interleaving.  Every line has its
    own cache key here, so the first run learns and the second run hits (CompressionBatch.java:47-52 warm start)."""
    params = pkg.CompressionParams(0.5, 0, 320, 320, 30000)
    items = []
    for i in range(24):
        if i % 4 == 3:
            items.append(batch.WorkItem(i, "png", synth.natural_image(870 + i, 700 + 8 * i, 500, channels=4), 10 ** 6))
        elif i % 7 == 6:
            items.append(batch.WorkItem(i, "gif", None, 1234))
        else:
            items.append(batch.WorkItem(i, "jpg", synth.natural_image(870 + i, 640 + 16 * (i % 5), 480), 102400 * (i + 3)))
    cache = {}
    for run in range(2):
        outputs = {}
        lock = threading.Lock()

        def on_output(it, out):
            with lock:
                outputs[it.index] = bytes(out) if it.kind == "jpg" else np.array(out)

        before = dict(cache)
        c = batch.run_shard(engine, items, params, cache, threads=3, inflight_per_thread=2, on_output=on_output)
        assert c.counts[batch.COMPRESSED_SUCCESS] == sum(1 for it in items if it.kind != "gif")
        assert c.counts[batch.FAILED_UNSUPPORTED_FORMAT] == sum(1 for it in items if it.kind == "gif")
        total = 0
        for it in items:
            if it.kind == "jpg":
                key = pkg.create_key(it.pixels.shape[1], it.pixels.shape[0], it.original_size)
                lp = before.get(key)
                rc, ref, rq, rs, uc, _ = oracle.jpg_fit(it.pixels, params.quality, params.target_max_size_bytes,
                                                        cached=(lp.quality, lp.scale) if lp else None)
                assert rc == 0 and outputs[it.index] == ref
                assert bool(uc) == (run == 1)
                assert cache[key] == LearnedParams(rq, rs) if run == 0 else cache[key] == lp
                total += len(ref)
            elif it.kind == "png":
                dw, dh, _ = pkg.png_target_dims(it.pixels.shape[1], it.pixels.shape[0], 320, 320)
                assert np.array_equal(outputs[it.index], oracle.resize(it.pixels, dw, dh, fmt=oracle.FMT_RGBA8))
                total += dw * dh * 4
        assert c.total_compressed_size == total + 1234 * sum(1 for it in items if it.kind == "gif")   # unsupported lines count (orig, orig)
