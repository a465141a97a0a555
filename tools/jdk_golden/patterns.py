"""Deterministic test images defined by 32-bit integer arithmetic only, so that
tools/jdk_golden/GoldenDump.java (on a JDK box) and the Python tests (here) build the
SAME pixels without exchanging files.  Nothing here is statistics-faithful; the
natural-statistics generator of the benchmark is image_compression_b200/synth.py.

    pattern(kind, w, h, channels) -> uint8 (h, w, channels), channel order R, G, B[, A]
"""
from __future__ import annotations

import numpy as np

KINDS = ("grad", "noise", "mix")
M32 = 0xFFFFFFFF


def _hash(x: np.ndarray, y: np.ndarray, c: int) -> np.ndarray:
    """GoldenDump.hash(): all operations mod 2^32, logical right shifts."""
    v = ((x * np.uint64(73856093)) ^ (y * np.uint64(19349663)) ^ np.uint64((c * 83492791) & M32)) & np.uint64(M32)
    v = ((v ^ (v >> np.uint64(13))) * np.uint64(0x5BD1E995)) & np.uint64(M32)
    v = v ^ (v >> np.uint64(15))
    return (v & np.uint64(0xFF)).astype(np.int32)


def pattern(kind: str, w: int, h: int, channels: int = 3) -> np.ndarray:
    if kind not in KINDS:
        raise ValueError(kind)
    x = np.arange(w, dtype=np.uint64)[None, :].repeat(h, 0)
    y = np.arange(h, dtype=np.uint64)[:, None].repeat(w, 1)
    xi, yi = x.astype(np.int64), y.astype(np.int64)
    out = np.zeros((h, w, channels), np.uint8)
    for c in range(channels):
        if c == 0:
            g = xi * 255 // max(w - 1, 1)
        elif c == 1:
            g = yi * 255 // max(h - 1, 1)
        elif c == 2:
            g = (xi + yi) * 255 // max(w + h - 2, 1)
        else:                                   # alpha: opaque core, soft left/top ramps, a transparent corner
            g = np.minimum(255, np.minimum(xi * 8, yi * 8))
        n = _hash(x, y, c).astype(np.int64)
        if kind == "grad":
            v = g
        elif kind == "noise":
            v = n
        else:
            v = g + (n & 31) - 16 + ((((xi >> 3) + (yi >> 3)) & 1) * 24)
        out[:, :, c] = np.clip(v, 0, 255).astype(np.uint8)
    return out


# the vectors GoldenDump writes and tests/test_jdk_golden.py reads (keep the two lists in step)
JPEG_CASES = [("grad", 8, 8), ("mix", 16, 16), ("mix", 17, 17), ("noise", 40, 56), ("mix", 333, 251), ("mix", 1000, 24)]
RESIZE_CASES = [("mix", 800, 600, 3, 0.125), ("mix", 333, 251, 3, 0.85), ("mix", 333, 251, 3, 0.5),
                ("mix", 256, 192, 4, 0.5), ("mix", 256, 192, 4, 0.3), ("noise", 64, 48, 4, 0.7225),
                ("mix", 3840, 2160, 4, 0.5), ("mix", 7680, 4320, 4, 0.25)]
FIT_CASES = [("mix", 1200, 900, 0.25, 60000), ("mix", 1200, 900, 1.0, 20000), ("mix", 1200, 900, 0.25, 3000),
             ("noise", 640, 480, 0.25, 20000)]
