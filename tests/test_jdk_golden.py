"""Parity against the REAL reference stack, when its vectors exist.

tools/jdk_golden/GoldenDump.java, run on any JDK 21 box (with the reference's shadow jar on the classpath
for the `fit` section), writes tests/golden/jdk_golden.json: javax.imageio JPEG bytes for every candidate
quality k/128 as core/ImageCompressionJpg.java:136-147 produces them, Graphics2D bilinear pixels as
tools/ImageTools.java:18-24 produces them, and the outcome of the reference's own
compressJpgWithTargetSize (core/ImageCompressionJpg.java:77-122).  This build image has no JVM, so the
file is absent here and these tests report "parity unpinned" as an expected failure; dropping the file
in flips them to real checks of the oracle (CPU suite) and of the CUDA path (-m gpu) with no code change.
Inputs are rebuilt bit for bit from the integer patterns of tools/jdk_golden/patterns.py.
"""
import hashlib
import json
import os
import struct
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tools", "jdk_golden"))
import patterns  # noqa: E402

GOLDEN = os.path.join(ROOT, "tests", "golden", "jdk_golden.json")
UNPINNED = ("parity unpinned: tests/golden/jdk_golden.json is absent (no JVM in the build image); "
            "run tools/jdk_golden/GoldenDump.java on a JDK 21 box to pin the oracle against javax.imageio / Graphics2D")


def _golden():
    if not os.path.exists(GOLDEN):
        pytest.xfail(UNPINNED)
    with open(GOLDEN) as f:
        return json.load(f)


def _sha(b) -> str:
    return hashlib.sha256(bytes(b)).hexdigest()


def test_patterns_are_stable():
    """The generator both sides share must never drift: pinned digests of a few patterns."""
    got = {k: _sha(patterns.pattern(k, 17, 9, 4).tobytes())[:12] for k in patterns.KINDS}
    assert got == {"grad": "344cca81ccb2", "noise": "40a21b17cd96", "mix": "ff515f0f01d0"}
    assert patterns.pattern("mix", 5, 4, 3).shape == (4, 5, 3)


def test_case_lists_match_the_java_side():
    """GoldenDump.java carries the same case tables as patterns.py."""
    text = open(os.path.join(ROOT, "tools", "jdk_golden", "GoldenDump.java")).read()
    for kind, w, h in patterns.JPEG_CASES:
        assert '{"%s", "%d", "%d"}' % (kind, w, h) in text
    for kind, w, h, c, s in patterns.RESIZE_CASES:
        assert '{"%s", "%d", "%d", "%d", "%s"}' % (kind, w, h, c, repr(s)) in text
    for kind, w, h, q, t in patterns.FIT_CASES:
        assert '{"%s", "%d", "%d", "%s", "%d"}' % (kind, w, h, repr(q), t) in text


def _check_jpeg(encode, g):
    for rec in g["jpeg"]:
        img = patterns.pattern(rec["kind"], rec["w"], rec["h"], 3)
        q = float(np.float32(rec["k"]) / np.float32(128.0))
        data = encode(img, q)
        head = bytes.fromhex(rec["head"])
        assert data[: len(head)] == head[: len(data)], ("header/tables differ", rec["kind"], rec["w"], rec["h"], rec["k"])
        assert (len(data), _sha(data)) == (rec["len"], rec["sha256"]), (rec["kind"], rec["w"], rec["h"], rec["k"])


def _check_resize(resize, g, max_pixels):
    for rec in g["resize"]:
        if rec["w"] * rec["h"] > max_pixels:
            continue
        img = patterns.pattern(rec["kind"], rec["w"], rec["h"], rec["channels"])
        out = resize(img, rec["dw"], rec["dh"], rec["channels"])
        assert out.shape == (rec["dh"], rec["dw"], rec["channels"])
        if rec.get("pixels"):
            ref = np.frombuffer(bytes.fromhex(rec["pixels"]), np.uint8).reshape(out.shape)
            assert np.abs(out.astype(int) - ref.astype(int)).max() <= 1, "outside the stated +-1 LSB"
        assert _sha(out.tobytes()) == rec["sha256"], "within tolerance but not bit-exact: correct the [JDK-mem] loop"


def _check_fit(fit, g):
    if not g["fit"]:
        pytest.xfail("jdk_golden.json was written without the reference jar on the classpath: search loop unpinned")
    for rec in g["fit"]:
        img = patterns.pattern(rec["kind"], rec["w"], rec["h"], 3)
        status, data, best_q, scale = fit(img, float(rec["init_q"]), rec["target"])
        assert (status == 0) == rec["ok"]
        if rec["ok"]:
            assert (len(data), _sha(data)) == (rec["len"], rec["sha256"])
            lq = struct.unpack("<f", struct.pack("<i", rec["learned"]["quality_bits"]))[0]
            ls = struct.unpack("<d", struct.pack("<q", rec["learned"]["scale_bits"]))[0]
            assert (best_q, scale) == (lq, ls)


def _check_decode(decode, g):
    if "decode" not in g:
        pytest.xfail("jdk_golden.json predates the decode section: decoded-pixel parity unpinned")
    for rec in g["decode"]:
        data = bytes.fromhex(rec["file"])
        px = decode(data)
        assert px.shape == (rec["h"], rec["w"], 3)
        assert _sha(px.tobytes()) == rec["sha256"], ("decoded raster differs from ImageIO's", rec["kind"], rec["w"], rec["h"], rec["k"])


# ---- the oracle against the JDK (CPU suite) -------------------------------------------------------
def test_oracle_decode_equals_imageio(oracle):
    """core/ImageCompression.java:119-155: the reader the reference gets from ImageIO, on the JDK's own files"""
    _check_decode(lambda data: oracle.jpeg_decode(data), _golden())


def test_oracle_jpeg_bytes_equal_imageio(oracle):
    _check_jpeg(lambda img, q: oracle.encode(img, q), _golden())


def test_oracle_resize_equals_graphics2d(oracle):
    fm = {3: oracle.FMT_RGB8, 4: oracle.FMT_RGBA8}
    _check_resize(lambda img, dw, dh, c: oracle.resize(img, dw, dh, fmt=fm[c]), _golden(), 4000 * 3000)


def test_oracle_fit_equals_the_reference_loop(oracle):
    def fit(img, q, target):
        rc, data, bq, sc, _, _ = oracle.jpg_fit(img, q, target)
        return rc, data, bq, sc
    _check_fit(fit, _golden())


def test_writer_is_the_one_the_survey_assumes():
    """TwelveMonkeys' writer SPI sits first and delegates to the JDK writer [JDK-mem]."""
    g = _golden()
    assert "JPEGImageWriter" in g["env"]["writer_class"], g["env"]


# ---- the CUDA path against the JDK (-m gpu) --------------------------------------------------------
@pytest.mark.gpu
def test_gpu_jpeg_bytes_equal_imageio(engine):
    _check_jpeg(lambda img, q: engine.compress_jpg_to_stream(img, q), _golden())


@pytest.mark.gpu
def test_gpu_resize_equals_graphics2d(engine):
    _check_resize(lambda img, dw, dh, c: engine.resize_to(img, dw, dh), _golden(), 1 << 40)


@pytest.mark.gpu
def test_gpu_fit_equals_the_reference_loop(engine):
    def fit(img, q, target):
        r = engine.jpg_fit(img, q, target)
        return r.status, r.data, r.best_q, r.scale
    _check_fit(fit, _golden())


@pytest.mark.gpu
def test_gpu_decode_equals_imageio(engine):
    _check_decode(lambda data: engine.jpg_decode(data), _golden())
