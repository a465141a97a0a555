"""GPU suite, SURVEY section 8 row f2: the JPEG reader on the GPU (csrc/k_decode.cuh) against the decode oracle
and the committed libjpeg-turbo fixtures -- bit-exact pixels -- and icb_jpg_file_fit (decode + decode rule + fit,
core/ImageCompression.java:119-155 + core/ImageCompressionJpg.java:77-122) against the oracle run on the decoded
pixels.  Everything goes through the C ABI."""
import io
import os

import numpy as np
import pytest

import image_compression_b200 as pkg
from image_compression_b200 import api, synth

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))
Z = np.load(os.path.join(HERE, "golden", "jpeg_decode.npz"))
N = len([k for k in Z.files if k.startswith("file_")])


@pytest.mark.parametrize("i", range(N))
def test_decode_equals_the_libjpeg_turbo_fixture(engine, oracle, i):
    data, want = Z[f"file_{i}"], Z[f"pixels_{i}"]
    got = engine.jpg_decode(data)
    assert np.array_equal(got, want)
    assert np.array_equal(engine.jpg_decode(data, api.FMT_BGR8), want[:, :, ::-1])


@pytest.mark.parametrize("w,h,q,kind", [(1000, 750, 0.9, "natural"), (4000, 3000, 0.92, "natural"), (1531, 1023, 1.0, "noise"),
                                        (640, 481, 0.05, "natural"), (2049, 33, 0.98, "noise"), (6000, 4000, 0.75, "natural")])
def test_decode_equals_the_oracle_on_jdk_style_files(engine, oracle, w, h, q, kind):
    """files as the JDK writer lays them out (oracle.encode: standard tables, stuffed 0xFF bytes at q = 1 on noise)"""
    img = (np.random.default_rng(w).integers(0, 256, (h, w, 3), dtype=np.uint8) if kind == "noise"
           else synth.natural_image(500 + w, w, h))
    data = oracle.encode(img, q)
    want = oracle.jpeg_decode(data)
    try:
        got = engine.jpg_decode(data)
    except pkg.IcbError as e:
        # noise at q = 1 has no end-of-block codes at all (every block runs to coefficient 63), so a decoder that
        # starts inside a block never finds the zigzag index again: the one kind of stream the parallel decoder
        # may give back to the host decoder -- as a refusal, never as wrong pixels
        assert kind == "noise" and q == 1.0 and e.code == api.ICB_ERR_UNSUPPORTED
        return
    assert np.array_equal(got, want)


def test_file_fit_equals_fit_on_the_decoded_pixels(engine, oracle):
    img = synth.natural_image(610, 1600, 1200)
    data = oracle.encode(img, 0.95)
    px = oracle.jpeg_decode(data)
    for q, target in ((0.25, 60000), (0.5, 200000), (1.0, 30000)):
        rc, ref, rq, rs, _, _ = oracle.jpg_fit(px, q, target)
        r = engine.jpg_file_fit(data, q, target)
        assert (r.status, r.best_q, r.scale) == (rc, rq, rs)
        if rc == 0:
            assert r.data == ref
    # cached parameters ride along like in icb_jpg_fit
    c = pkg.LearnedParams(0.2, 1.0)
    rc, ref, rq, rs, used, _ = oracle.jpg_fit(px, 0.25, 100000, cached=(0.2, 1.0))
    r = engine.jpg_file_fit(data, 0.25, 100000, cached=c)
    assert (r.status, r.best_q, r.scale, r.used_cache) == (rc, rq, rs, bool(used)) and r.data == ref


def test_decode_rule_subsamples_large_files_on_the_gpu(engine, oracle):
    """core/ImageCompression.java:140-153: longer side 8200 -> setSourceSubsampling(2, 2, 0, 0)"""
    img = synth.natural_image(620, 8200, 300)
    data = oracle.encode(img, 0.9)
    px = oracle.subsample(oracle.jpeg_decode(data), 2)
    rc, ref, rq, rs, _, _ = oracle.jpg_fit(px, 0.25, 50000)
    r = engine.jpg_file_fit(data, 0.25, 50000)
    assert (r.status, r.best_q, r.scale) == (rc, rq, rs) and r.data == ref


@pytest.mark.parametrize("name", ["progressive", "gray"])
def test_other_kinds_of_jpeg_are_refused(engine, name):
    with pytest.raises(pkg.IcbError) as ei:
        engine.jpg_decode(Z["unsup_" + name])
    assert ei.value.code == api.ICB_ERR_UNSUPPORTED


def test_broken_files_are_errors_not_crashes(engine):
    data = Z["file_2"]
    for cut in (0, 3, 100, 700, len(data) // 2, len(data) - 40):
        with pytest.raises(pkg.IcbError):
            engine.jpg_decode(np.ascontiguousarray(data[:cut]))
    # a good file still decodes afterwards
    assert np.array_equal(engine.jpg_decode(data), Z["pixels_2"])


def test_file_items_in_the_native_executor(engine, oracle):
    imgs = [synth.natural_image(630 + i, 800 + 16 * i, 600) for i in range(6)]
    files = [oracle.encode(im, 0.9) for im in imgs]
    outs = [np.empty(40000, np.uint8) for _ in files]
    items = engine.make_items([dict(kind="jpg_file_fit", file=f, init_q=0.25, target=40000, out=o)
                               for f, o in zip(files, outs)])
    engine.run_list(items, threads=2, inflight=2)
    for f, o, r, (dw, dh) in zip(files, outs, items.results(), items.dims()):
        px = oracle.jpeg_decode(f)
        rc, ref, rq, rs, _, _ = oracle.jpg_fit(px, 0.25, 40000)
        assert (r.status, r.best_q, r.scale) == (rc, rq, rs) and (dw, dh) == (px.shape[1], px.shape[0])
        if rc == 0:
            assert bytes(o[: r.out_len]) == ref


def test_run_shard_takes_file_lines_and_falls_back_to_pixels(engine, oracle):
    """batch.run_shard (the CompressionBatch mirror): JPG lines that carry their file go up as files; a line whose file
    is of another kind (here: progressive) runs on its host-decoded pixels; results and learned cache equal the oracle's."""
    from image_compression_b200 import batch
    params = pkg.CompressionParams(0.25, 0, 100, 100, 30000)
    items, expect = [], {}
    for i in range(8):
        img = synth.natural_image(700 + i, 800 + 32 * i, 600)
        data = oracle.encode(img, 0.9)
        px = oracle.jpeg_decode(data)
        items.append(batch.WorkItem(i, "jpg", None, 102400 * (i + 2), file_bytes=data))
        expect[i] = (px, 102400 * (i + 2))
    px444 = synth.natural_image(720, 640, 480)
    items.append(batch.WorkItem(8, "jpg", px444, 999999, file_bytes=Z["unsup_progressive"].tobytes()))
    expect[8] = (px444, 999999)
    cache, outputs = {}, {}
    c = batch.run_shard(engine, items, params, cache, threads=2, inflight_per_thread=2,
                        on_output=lambda it, out: outputs.__setitem__(it.index, bytes(out)))
    assert c.counts[batch.COMPRESSED_SUCCESS] == len(items)
    for i, (px, size) in expect.items():
        rc, ref, rq, rs, _, _ = oracle.jpg_fit(px, params.quality, params.target_max_size_bytes)
        assert rc == 0 and outputs[i] == ref
        assert cache[pkg.create_key(px.shape[1], px.shape[0], size)] == pkg.LearnedParams(rq, rs)


@pytest.mark.parametrize("sub,w,h", [(1, 1203, 901), (0, 1203, 901), (1, 4000, 3000), (0, 2048, 1536)])
def test_422_and_444_files_decode_like_the_oracle(engine, oracle, sub, w, h):
    """4:2:2 (h2v1 fancy upsampling) and 4:4:4 files at sizes the committed fixtures do not reach, written by Pillow here;
    the oracle (pinned against libjpeg-turbo on the small fixtures of both kinds) is the checker, and file_fit rides along."""
    Image = pytest.importorskip("PIL.Image")
    img = synth.natural_image(800 + w + sub, w, h)
    bio = io.BytesIO()
    Image.fromarray(img).save(bio, "JPEG", quality=90, subsampling=sub)
    data = bio.getvalue()
    want = oracle.jpeg_decode(data)
    assert np.array_equal(engine.jpg_decode(data), want)
    if w < 2000:
        rc, ref, rq, rs, _, _ = oracle.jpg_fit(want, 0.25, 60000)
        r = engine.jpg_file_fit(data, 0.25, 60000)
        assert (r.status, r.best_q, r.scale) == (rc, rq, rs) and r.data == ref


@pytest.mark.parametrize("w,h,sub,kw", [(4000, 3000, 2, dict(restart_marker_rows=1)), (1531, 1023, 2, dict(restart_marker_blocks=4)),
                                        (1203, 901, 1, dict(restart_marker_blocks=1)), (2048, 1536, 0, dict(restart_marker_rows=3))])
def test_files_with_restart_intervals(engine, oracle, w, h, sub, kw):
    """DRI / RSTn: markers dropped with the stuffing, every interval start a known entry point of the parallel decoder,
    DC predictions starting again per interval -- at sizes beyond the committed fixtures (files written by Pillow here)"""
    Image = pytest.importorskip("PIL.Image")
    img = synth.natural_image(900 + w, w, h)
    bio = io.BytesIO()
    Image.fromarray(img).save(bio, "JPEG", quality=88, subsampling=sub, **kw)
    data = bio.getvalue()
    assert b"\xff\xdd" in data
    want = oracle.jpeg_decode(data)
    assert np.array_equal(engine.jpg_decode(data), want)


@pytest.mark.parametrize("variant", ["extra_segments", "merged_tables", "fill_and_tail"])
def test_rewrapped_files_decode_alike(engine, variant):
    """what camera files look like: Exif / comment segments, one DHT / DQT segment for all tables, fill bytes in front
    of markers, bytes behind EOI (tests/jpeg_mangle.py) -- same pixels as the plain file"""
    import sys
    sys.path.insert(0, HERE)
    import jpeg_mangle
    for i in (2, 13, 19, 24):
        data = jpeg_mangle.VARIANTS[variant](Z[f"file_{i}"].tobytes())
        assert np.array_equal(engine.jpg_decode(np.frombuffer(data, np.uint8)), Z[f"pixels_{i}"]), (variant, i)
