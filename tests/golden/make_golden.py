#!/usr/bin/env python
"""Generates the golden fixtures under tests/golden/ (run once, in the build container).

The reference (PolloChang/image-compression) holds NO golden vector for the JPG path
and only output-dimension assertions for PNG (src/test/.../ImageCompressionPngTest.java
:38-88), and its arithmetic lives in the JDK (libjavajpeg = IJG 6b, Java2D loops), which
cannot run here (no JVM).  The fixtures therefore come from INDEPENDENT implementations
of the same published algorithms that exist in this container -- never from oracle/ or
from the CUDA library:

  jpeg_pillow.npz   Pillow (libjpeg-turbo, bit-exact with IJG 6b for baseline/islow)
                    fed the quantisation tables explicitly; APP0 version byte 12
                    patched 01 -> 02 (the one known JDK difference, [JDK-mem]).
  qtables.npz       numpy float32 evaluation of JDK JPEG.convertToLinearQuality +
                    JPEGQTable.getScaledInstance for every candidate k/128, and the
                    IJG integer formula for the alternative profile.
  search.json       the reference's float32 bisection
                    (core/ImageCompressionJpg.java:158-200) driven in numpy float32
                    over Pillow-encoded sizes: visited midpoints, sizes, result.
  resize_cv2.npz    OpenCV INTER_LINEAR (pixel-centre 2x2 bilinear, no area
                    averaging) -- pins the resize within the +-1 LSB tolerance that
                    BASELINE.json states; dims rule from tools/ImageTools.java:8-9
                    and the reference's own PNG tests (800x600 -> 100x75).

    python tests/golden/make_golden.py
"""
from __future__ import annotations

import hashlib
import io
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

BASE_LUM = np.array([
    16, 11, 10, 16, 24, 40, 51, 61, 12, 12, 14, 19, 26, 58, 60, 55,
    14, 13, 16, 24, 40, 57, 69, 56, 14, 17, 22, 29, 51, 87, 80, 62,
    18, 22, 37, 56, 68, 109, 103, 77, 24, 35, 55, 64, 81, 104, 113, 92,
    49, 64, 78, 87, 103, 121, 120, 101, 72, 92, 95, 98, 112, 100, 103, 99], np.int32)
BASE_CHR = np.array([
    17, 18, 24, 47, 99, 99, 99, 99, 18, 21, 26, 66, 99, 99, 99, 99,
    24, 26, 56, 99, 99, 99, 99, 99, 47, 66, 99, 99, 99, 99, 99, 99] + [99] * 32, np.int32)


def jdk_tables(q):
    """JPEG.convertToLinearQuality + JPEGQTable.getScaledInstance(scale, forceBaseline) in float32."""
    f = np.float32
    q = f(q)
    if q <= f(0):
        q = f(0.01)
    if q > f(1):
        q = f(1)
    s = f(0.5) / q if q < f(0.5) else f(2) - q * f(2)
    out = []
    for base in (BASE_LUM, BASE_CHR):
        v = (base.astype(np.float32) * f(s) + f(0.5)).astype(np.float32)
        out.append(np.clip(v.astype(np.int32), 1, 255).astype(np.uint8))
    return out


def ijg_tables(q):
    quality = int(np.float32(q) * np.float32(100) + np.float32(0.5))
    quality = min(max(quality, 1), 100)
    scale = 5000 // quality if quality < 50 else 200 - quality * 2
    return [np.clip((b.astype(np.int64) * scale + 50) // 100, 1, 255).astype(np.uint8) for b in (BASE_LUM, BASE_CHR)]


def pillow_encode(img, lum, chr_, jfif_minor=2):
    from PIL import Image
    b = io.BytesIO()
    Image.fromarray(img).save(b, format="JPEG", qtables=[[int(x) for x in lum], [int(x) for x in chr_]],
                              subsampling=2, optimize=False)
    v = bytearray(b.getvalue())
    assert v[6:11] == b"JFIF\0" and v[11] == 1
    v[12] = jfif_minor
    return bytes(v)


def small_images():
    """Known-answer inputs stored verbatim (not re-generated at test time)."""
    from image_compression_b200 import synth
    out = {}
    for (w, h) in [(1, 1), (8, 8), (16, 16), (17, 17), (40, 56), (33, 15), (15, 33), (257, 49)]:
        out[f"nat_{w}x{h}"] = synth.natural_image(100 + w + 3 * h, w, h)
    out["noise_64x48"] = synth.noise_image(7, 64, 48)          # 0xFF byte-stuffing stress at q=1
    out["noise_120x24"] = synth.noise_image(8, 120, 24)
    out["checker_50x35"] = synth.pattern_image("checker", 50, 35)
    out["solid_24x24"] = synth.pattern_image("solid", 24, 24)
    out["gradient_96x40"] = synth.pattern_image("gradient", 96, 40)
    return out


def java_search(size_of, target, initial_q):
    """core/ImageCompressionJpg.java:158-200 in float32."""
    f = np.float32
    lo, hi, best = f(0), f(initial_q), f(-1)
    trace = []
    for _ in range(8):
        mid = (lo + hi) / f(2)
        if mid < f(0.01):
            break
        size = size_of(float(mid))
        trace.append([float(mid), int(size)])
        if size <= target:
            best = mid
            lo = mid
        else:
            hi = mid
        if (hi - lo) < f(0.01):
            break
    return float(best), trace


def main():
    import cv2
    from image_compression_b200 import synth

    # ---- quantisation tables of every candidate
    ks = np.arange(1, 129)
    jd = np.stack([np.stack(jdk_tables(k / 128.0)) for k in ks])
    ij = np.stack([np.stack(ijg_tables(k / 128.0)) for k in ks])
    extra_q = np.array([0.0, 0.005, 0.01, 0.0100001, 0.3, 0.49999, 0.5, 0.75, 0.999, 1.0, 1.5], np.float32)
    jd_x = np.stack([np.stack(jdk_tables(q)) for q in extra_q])
    ij_x = np.stack([np.stack(ijg_tables(q)) for q in extra_q])
    np.savez_compressed(os.path.join(HERE, "qtables.npz"), k=ks, jdk=jd, ijg=ij, extra_q=extra_q, jdk_extra=jd_x,
                        ijg_extra=ij_x)

    # ---- JFIF bytes from libjpeg-turbo
    imgs = small_images()
    qs = [0.015625, 0.1, 0.2421875, 0.5, 0.75, 0.9921875, 1.0]
    pack = {}
    index = []
    for name, img in imgs.items():
        pack["img_" + name] = img
        for q in qs:
            lum, chr_ = jdk_tables(q)
            data = pillow_encode(img, lum, chr_)
            key = f"{name}@{q}"
            pack["jpg_" + key] = np.frombuffer(data, np.uint8)
            index.append({"image": name, "q": q, "len": len(data), "sha256": hashlib.sha256(data).hexdigest()})
    np.savez_compressed(os.path.join(HERE, "jpeg_pillow.npz"), **pack)

    # ---- larger images: hashes only (regenerated from the seeded generator at test time)
    big = []
    for (idx, w, h) in [(3, 333, 251), (4, 640, 480), (5, 1023, 767), (6, 4000, 24), (9, 24, 1000)]:
        img = synth.natural_image(idx, w, h)
        pix = hashlib.sha256(img.tobytes()).hexdigest()
        for q in (0.05, 0.2421875, 0.6, 0.95):
            lum, chr_ = jdk_tables(q)
            data = pillow_encode(img, lum, chr_)
            big.append({"index": idx, "w": w, "h": h, "q": q, "pixels_sha256": pix, "len": len(data),
                        "sha256": hashlib.sha256(data).hexdigest()})

    # ---- search traces (Pillow sizes + float32 bisection)
    traces = []
    img = synth.natural_image(21, 640, 480)
    pix = hashlib.sha256(img.tobytes()).hexdigest()
    cache = {}

    def size_of(q):
        if q not in cache:
            cache[q] = len(pillow_encode(img, *jdk_tables(q)))
        return cache[q]

    for init_q in (0.25, 0.5, 0.75, 1.0):
        full = size_of(float(np.float32(init_q)))
        for frac in (0.001, 0.02, 0.3, 0.6, 0.9, 1.2, 5.0):
            target = int(full * frac)
            best, tr = java_search(size_of, target, init_q)
            traces.append({"init_q": init_q, "target": target, "best": best, "trace": tr})
    # a target that lands exactly on a candidate's size, and one byte under it
    s = size_of(0.125)
    for target in (s, s - 1):
        best, tr = java_search(size_of, target, 0.25)
        traces.append({"init_q": 0.25, "target": target, "best": best, "trace": tr})
    json.dump({"image": {"index": 21, "w": 640, "h": 480, "pixels_sha256": pix}, "cases": traces, "big": big},
              open(os.path.join(HERE, "search.json"), "w"), indent=0)

    # ---- resize
    rz = {}
    meta = []
    # 800x600 -> 100x75 is the reference's own PNG test case (ImageCompressionPngTest.java:54-70);
    # the source is regenerated from its seed at test time (pixel hash in the meta), only outputs are stored
    src = synth.natural_image(31, 800, 600)
    rz["hash_800x600"] = np.array(hashlib.sha256(src.tobytes()).hexdigest())
    for (dw, dh) in [(100, 75), (200, 150), (187, 131), (1, 1)]:
        rz[f"dst_800x600_{dw}x{dh}"] = cv2.resize(src, (dw, dh), interpolation=cv2.INTER_LINEAR)
        meta.append(["seed31_800x600", dw, dh])
    src2 = synth.natural_image(32, 101, 77)
    rz["src_101x77"] = src2
    for (dw, dh) in [(85, 65), (50, 38), (10, 7)]:
        rz[f"dst_101x77_{dw}x{dh}"] = cv2.resize(src2, (dw, dh), interpolation=cv2.INTER_LINEAR)
        meta.append(["src_101x77", dw, dh])
    rz["meta"] = np.array(json.dumps(meta))
    np.savez_compressed(os.path.join(HERE, "resize_cv2.npz"), **rz)
    for f in sorted(os.listdir(HERE)):
        print(f, os.path.getsize(os.path.join(HERE, f)))


if __name__ == "__main__":
    main()
