"""CPU suite, part 1: pins the oracle (oracle/, test infrastructure) against the golden
fixtures in tests/golden/, which were produced by INDEPENDENT implementations
(Pillow/libjpeg-turbo, numpy float32, OpenCV) -- see tests/golden/make_golden.py.
The reference holds no JPG vectors of its own."""
import hashlib
import json
import os

import numpy as np
import pytest

from image_compression_b200 import synth

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


@pytest.fixture(scope="module")
def jpeg_gold():
    return np.load(os.path.join(GOLD, "jpeg_pillow.npz"))


@pytest.fixture(scope="module")
def search_gold():
    return json.load(open(os.path.join(GOLD, "search.json")))


def test_qtables_every_candidate(oracle):
    g = np.load(os.path.join(GOLD, "qtables.npz"))
    for i, k in enumerate(g["k"]):
        lum, chr_ = oracle.qtables(float(k) / 128.0, oracle.JDK)
        assert np.array_equal(lum, g["jdk"][i, 0]) and np.array_equal(chr_, g["jdk"][i, 1]), k
        lum, chr_ = oracle.qtables(float(k) / 128.0, oracle.IJG)
        assert np.array_equal(lum, g["ijg"][i, 0]) and np.array_equal(chr_, g["ijg"][i, 1]), k
    for i, q in enumerate(g["extra_q"]):
        lum, chr_ = oracle.qtables(float(q), oracle.JDK)
        assert np.array_equal(lum, g["jdk_extra"][i, 0]) and np.array_equal(chr_, g["jdk_extra"][i, 1]), q
        lum, chr_ = oracle.qtables(float(q), oracle.IJG)
        assert np.array_equal(lum, g["ijg_extra"][i, 0]) and np.array_equal(chr_, g["ijg_extra"][i, 1]), q


def test_candidate_extremes(oracle):
    # k=2 (scale 32) clamps every entry to 255; k=127 (scale 1/64) leaves only 1s and 2s
    lum, chr_ = oracle.qtables(2 / 128.0)
    assert lum.min() == 255 and chr_.min() == 255
    lum, chr_ = oracle.qtables(127 / 128.0)
    assert lum.max() == 2 and chr_.max() == 2 and lum.min() == 1


def test_jfif_bytes_small_images(oracle, jpeg_gold):
    keys = [k for k in jpeg_gold.files if k.startswith("jpg_")]
    assert len(keys) >= 90
    for k in keys:
        name, q = k[4:].split("@")
        img = jpeg_gold["img_" + name]
        got = oracle.encode(img, float(q))
        assert got == jpeg_gold[k].tobytes(), k


def test_jfif_header_layout(oracle, jpeg_gold):
    """a2.8: 623 header bytes + EOI, marker order SOI APP0 DQT DQT SOF0 DHTx4 SOS."""
    img = jpeg_gold["img_nat_17x17"]
    d = oracle.encode(img, 0.5)
    assert d[:4] == b"\xff\xd8\xff\xe0" and d[6:11] == b"JFIF\0" and d[11:13] == b"\x01\x02"
    assert d[20:22] == b"\xff\xdb" and d[89:91] == b"\xff\xdb" and d[158:160] == b"\xff\xc0"
    assert d[163:167] == bytes([0, 17, 0, 17])
    assert d[167:177] == bytes([3, 1, 0x22, 0, 2, 0x11, 1, 3, 0x11, 1])
    assert d[609:623] == bytes([0xff, 0xda, 0, 12, 3, 1, 0, 2, 0x11, 3, 0x11, 0, 0x3f, 0])
    assert d[-2:] == b"\xff\xd9"


def test_jfif_bytes_larger_images_by_hash(oracle, search_gold):
    cache = {}
    for e in search_gold["big"]:
        key = (e["index"], e["w"], e["h"])
        if key not in cache:
            cache[key] = synth.natural_image(*key)
            assert hashlib.sha256(cache[key].tobytes()).hexdigest() == e["pixels_sha256"], \
                "synthetic generator changed: regenerate tests/golden"
        got = oracle.encode(cache[key], e["q"])
        assert len(got) == e["len"]
        assert hashlib.sha256(got).hexdigest() == e["sha256"]


def test_search_traces(oracle, search_gold):
    m = search_gold["image"]
    img = synth.natural_image(m["index"], m["w"], m["h"])
    assert hashlib.sha256(img.tobytes()).hexdigest() == m["pixels_sha256"]
    for c in search_gold["cases"]:
        best, trace = oracle.find_best_quality(img, c["target"], c["init_q"])
        assert best == c["best"], c
        assert [[q, s] for q, s in trace] == c["trace"], c


def test_search_break_rules(oracle):
    """`mid < 0.01f` stops before encoding k=1 (:170); `hi - lo < 0.01f` stops after (:187)."""
    img = synth.natural_image(2, 64, 48)
    best, trace = oracle.find_best_quality(img, 1, 0.25)            # nothing fits
    assert best == -1.0
    assert [q for q, _ in trace] == [0.125, 0.0625, 0.03125, 0.015625]
    best, trace = oracle.find_best_quality(img, 10 ** 9, 0.25)     # everything fits
    assert [q for q, _ in trace] == [0.125, 0.1875, 0.21875, 0.234375, 0.2421875]
    assert best == 0.2421875
    best, trace = oracle.find_best_quality(img, 10 ** 9, 1.0)
    assert len(trace) == 7 and best == 127 / 128.0


def test_fit_scale_loop_and_cache(oracle):
    """:96-115 -- when even q=1/64 overshoots, the image is re-scaled by 0.85^n from the ORIGINAL."""
    img = synth.noise_image(3, 200, 150)
    smallest = oracle.encoded_size(img, 0.015625)
    rc, data, q, scale, used, n_enc = oracle.jpg_fit(img, 0.25, smallest - 1)
    assert rc == 0 and scale < 1.0 and abs(np.log(scale) / np.log(0.85) - round(np.log(scale) / np.log(0.85))) < 1e-9
    assert len(data) <= smallest - 1 and not used
    dw, dh = oracle.resize_dims(200, 150, scale)
    assert data == oracle.encode(oracle.resize(img, dw, dh), q)
    # nothing ever fits -> false (:120-121)
    rc, *_ = oracle.jpg_fit(img, 0.25, 600)
    assert rc == 1
    # cached params: hit (:228-231), and miss falling back to the full search (:86-88)
    t = oracle.encoded_size(img, 0.2) + 5
    rc, data, q, scale, used, n_enc = oracle.jpg_fit(img, 0.25, t, cached=(0.2, 1.0))
    assert (rc, q, scale, used, n_enc) == (0, np.float32(0.2), 1.0, 1, 1)
    rc, data, q, scale, used, n_enc = oracle.jpg_fit(img, 0.25, t, cached=(0.9, 1.0))
    assert rc == 0 and not used and n_enc > 2


def test_resize_dims_rule(oracle):
    # tools/ImageTools.java:8-9 and the reference's PNG tests (ImageCompressionPngTest.java:38-70)
    assert oracle.resize_dims(800, 600, 0.125) == (100, 75)
    assert oracle.resize_dims(4001, 3001, 0.85) == (3400, 2550)
    assert oracle.resize_dims(3, 3, 0.1) == (1, 1)
    assert oracle.png_target_dims(100, 100, 100, 100)[0] == 1                     # no resize -> false
    assert oracle.png_target_dims(800, 600, 100, 100)[:3] == (0, 100, 75)
    assert oracle.png_target_dims(3840, 2160, 1920, 1920)[:3] == (0, 1920, 1080)
    assert oracle.png_target_dims(7680, 4320, 1920, 1920)[:3] == (0, 1920, 1080)
    assert oracle.png_target_dims(50, 4000, 1920, 1920)[:3] == (0, 24, 1920)      # one dim small: still resized


def test_resize_pixels_within_one_lsb_of_opencv(oracle):
    g = np.load(os.path.join(GOLD, "resize_cv2.npz"))
    meta = json.loads(str(g["meta"]))
    big = synth.natural_image(31, 800, 600)
    assert hashlib.sha256(big.tobytes()).hexdigest() == str(g["hash_800x600"])
    for name, dw, dh in meta:
        src = big if name == "seed31_800x600" else g[name]
        tag = "800x600" if name == "seed31_800x600" else name[4:]
        ref = g[f"dst_{tag}_{dw}x{dh}"].astype(int)
        got = oracle.resize(src, dw, dh).astype(int)
        assert np.abs(got - ref).max() <= 1, (name, dw, dh)                       # BASELINE tolerance
        f64 = oracle.resize(src, dw, dh, f64=True).astype(int)
        assert np.abs(f64 - ref).max() <= 1, (name, dw, dh)


def test_resize_alpha_premultiplied_roundtrip(oracle):
    img = synth.natural_image(5, 90, 70, channels=4)
    opaque = img.copy()
    opaque[..., 3] = 255
    a = oracle.resize(opaque, 45, 35, fmt=oracle.FMT_RGBA8)
    b = oracle.resize(np.ascontiguousarray(opaque[..., :3]), 45, 35)
    assert np.array_equal(a[..., :3], b) and (a[..., 3] == 255).all()
    clear = img.copy()
    clear[..., 3] = 0
    assert not oracle.resize(clear, 45, 35, fmt=oracle.FMT_RGBA8).any()


def test_dummy_block_geometry(oracle):
    """a2.6: 4000x3000 has 375 luma block rows -> the last half-MCU row is dummy blocks (AC 0)."""
    g = oracle.geometry(4000, 3000)
    assert (g.mcu_cols, g.mcu_rows, g.yblk_h, g.nblocks) == (250, 188, 375, 282000)
    img = synth.natural_image(1, 40, 24)            # 3 luma block rows, 5 block cols -> dummies on both edges
    c = oracle.forward_coefficients(img).reshape(2, 3, 6, 64)
    assert not c[1, :, 2:4, 1:].any()               # bottom luma row of MCU row 1: dummy
    assert not c[:, 2, 1, 1:].any() and not c[0, 2, 3, 1:].any()   # right luma column of MCU col 2
    assert c[0, 2, 1, 0] == c[0, 2, 0, 0]           # dummy DC = DC of the block to its left
    assert c[0, 0, 0, 1:].any()


def test_subsample_rule_and_pixels(oracle):
    """core/ImageCompression.java:140-153 + ImageReadParam.setSourceSubsampling(s, s, 0, 0): the
    oracle against plain numpy slicing (an independent statement of "every s-th pixel")."""
    assert [oracle.subsample_factor(w, h) for w, h in [(4096, 10), (4097, 10), (8191, 8191), (8192, 1), (11548, 8660),
                                                        (16383, 5), (16384, 5), (70000, 3)]] == [1, 1, 1, 2, 2, 2, 4, 16]
    rng = np.random.default_rng(3)
    for (w, h, c) in [(17, 9, 3), (64, 48, 4), (101, 77, 3)]:
        img = rng.integers(0, 256, (h, w, c), dtype=np.uint8)
        for s in (1, 2, 3, 4, 8):
            assert np.array_equal(oracle.subsample(img, s), img[::s, ::s])
