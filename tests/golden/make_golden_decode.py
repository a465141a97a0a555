#!/usr/bin/env python
"""Generates tests/golden/jpeg_decode.npz (run once, in the build container): baseline 4:2:0 JPEG files written
by Pillow / libjpeg-turbo from the synthetic images, and the pixels Pillow / libjpeg-turbo decodes from them
(islow IDCT, fancy upsampling: bit-exact with the IJG-6b decoder the JDK's ImageIO reader wraps).  An
INDEPENDENT implementation: neither oracle/ nor the CUDA library takes part.

    python tests/golden/make_golden_decode.py
"""
import io
import os
import sys

import numpy as np
from PIL import Image

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
from image_compression_b200 import synth  # noqa: E402

# (seed, w, h, quality, kind, Pillow subsampling: 2 = 4:2:0, 1 = 4:2:2, 0 = 4:4:4)
CASES_OTHER = [(53, 97, 61, 92, "natural", 1), (54, 200, 133, 75, "natural", 1), (55, 16, 8, 95, "noise", 1), (56, 5, 7, 92, "natural", 1),
               (57, 3, 3, 92, "natural", 1), (58, 333, 100, 50, "natural", 1),
               (59, 97, 61, 92, "natural", 0), (60, 200, 133, 75, "natural", 0), (61, 8, 8, 98, "noise", 0), (62, 5, 7, 92, "natural", 0),
               (63, 131, 77, 30, "natural", 0)]

CASES = [  # (seed, w, h, quality, kind)
    (41, 64, 48, 92, "natural"), (42, 101, 75, 92, "natural"), (43, 200, 133, 75, "natural"),
    (44, 17, 9, 92, "natural"), (45, 333, 251, 50, "natural"), (46, 16, 16, 98, "noise"),
    (47, 130, 66, 100, "noise"), (48, 5, 3, 92, "natural"), (49, 3, 40, 92, "natural"),
    (50, 257, 31, 30, "natural"), (51, 96, 96, 10, "natural"), (52, 48, 160, 95, "noise"),
]


def make(seed, w, h, kind):
    if kind == "noise":
        return np.random.default_rng(seed).integers(0, 256, (h, w, 3), dtype=np.uint8)
    return synth.natural_image(seed, w, h)


def main():
    out = {}
    for i, (seed, w, h, q, kind) in enumerate(CASES):
        img = make(seed, w, h, kind)
        bio = io.BytesIO()
        Image.fromarray(img).save(bio, "JPEG", quality=q, subsampling=2, optimize=(i % 3 == 2))
        data = bio.getvalue()
        dec = np.asarray(Image.open(io.BytesIO(data)).convert("RGB"))
        assert dec.shape == (h, w, 3)
        out[f"file_{i}"] = np.frombuffer(data, np.uint8)
        out[f"pixels_{i}"] = dec
    for j, (seed, w, h, q, kind, sub) in enumerate(CASES_OTHER):
        img = make(seed, w, h, kind)
        bio = io.BytesIO()
        Image.fromarray(img).save(bio, "JPEG", quality=q, subsampling=sub, optimize=(j % 2 == 1))
        data = bio.getvalue()
        dec = np.asarray(Image.open(io.BytesIO(data)).convert("RGB"))
        out[f"file_{len(CASES) + j}"] = np.frombuffer(data, np.uint8)
        out[f"pixels_{len(CASES) + j}"] = dec
    # restart intervals (DRI + RSTn markers): whole MCU rows and a few MCUs, with and without a tail, three samplings
    base = len(CASES) + len(CASES_OTHER)
    for j, (seed, w, h, q, sub, kw) in enumerate([(70, 200, 133, 90, 2, dict(restart_marker_rows=1)), (71, 333, 251, 75, 2, dict(restart_marker_blocks=5)),
                                                   (72, 64, 48, 95, 2, dict(restart_marker_blocks=1)), (73, 97, 61, 92, 1, dict(restart_marker_rows=2)),
                                                   (74, 131, 77, 60, 0, dict(restart_marker_blocks=7)), (75, 130, 66, 100, 2, dict(restart_marker_blocks=3))]):
        img = make(seed, w, h, "noise" if seed == 75 else "natural")
        bio = io.BytesIO()
        Image.fromarray(img).save(bio, "JPEG", quality=q, subsampling=sub, **kw)
        data = bio.getvalue()
        assert b"\xff\xdd" in data
        dec = np.asarray(Image.open(io.BytesIO(data)).convert("RGB"))
        out[f"file_{base + j}"] = np.frombuffer(data, np.uint8)
        out[f"pixels_{base + j}"] = dec
    # files the decoder must refuse (the caller keeps the stock host decode): progressive, gray
    img = make(60, 48, 32, "natural")
    for name, kw in (("progressive", dict(progressive=True, subsampling=2)),):
        bio = io.BytesIO()
        Image.fromarray(img).save(bio, "JPEG", quality=90, **kw)
        out["unsup_" + name] = np.frombuffer(bio.getvalue(), np.uint8)
    bio = io.BytesIO()
    Image.fromarray(img[:, :, 0]).save(bio, "JPEG", quality=90)
    out["unsup_gray"] = np.frombuffer(bio.getvalue(), np.uint8)
    np.savez_compressed(os.path.join(HERE, "jpeg_decode.npz"), **out)
    print("wrote", len([k for k in out if k.startswith("file_")]), "cases")


if __name__ == "__main__":
    main()
