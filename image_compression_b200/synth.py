"""Deterministic synthetic "natural-statistics" images.

img = clip(G + T + N): G = sum of linear/radial colour gradients, T = 1/f^beta
texture by FFT spectral synthesis with correlated channels, N = Gaussian noise.
There is no network for datasets, so tests and bench.py use these.
"""
from __future__ import annotations

import numpy as np

SEED_BASE = 0xC0FFEE


def _texture(rng: np.random.Generator, h: int, w: int, beta: float) -> np.ndarray:
    """Zero-mean, unit-variance 1/f^beta field.  Synthesised at <= 1024 px on the long
    side and bilinearly upsampled (keeps 12-24 MP images to a couple of seconds)."""
    s = max(1, int(np.ceil(max(h, w) / 1024)))
    hh, ww = -(-h // s) + 1, -(-w // s) + 1
    fy = np.fft.fftfreq(hh)[:, None]
    fx = np.fft.rfftfreq(ww)[None, :]
    f = np.sqrt(fy * fy + fx * fx)
    f[0, 0] = 1.0
    amp = f ** (-beta / 2.0)
    amp[0, 0] = 0.0
    spec = amp * (rng.standard_normal(amp.shape) + 1j * rng.standard_normal(amp.shape))
    t = np.fft.irfft2(spec, s=(hh, ww)).astype(np.float32)
    t /= t.std() + 1e-9
    if s == 1:
        return t[:h, :w]
    ys = (np.arange(h, dtype=np.float32) / s)
    xs = (np.arange(w, dtype=np.float32) / s)
    y0 = np.minimum(ys.astype(np.int32), hh - 2)
    x0 = np.minimum(xs.astype(np.int32), ww - 2)
    fyv = (ys - y0)[:, None]
    fxv = (xs - x0)[None, :]
    rows = t[y0] * (1 - fyv) + t[y0 + 1] * fyv            # separable: rows first, then columns
    return rows[:, x0] * (1 - fxv) + rows[:, x0 + 1] * fxv


def natural_image(index: int, w: int, h: int, channels: int = 3) -> np.ndarray:
    """RGB8 (or RGBA8 with a smooth alpha) image number `index`, shape (h, w, channels)."""
    rng = np.random.default_rng(SEED_BASE + index)
    yy = np.linspace(0.0, 1.0, h, dtype=np.float32)[:, None]
    xx = np.linspace(0.0, 1.0, w, dtype=np.float32)[None, :]
    img = np.empty((h, w, 3), np.float32)
    ngrad = int(rng.integers(2, 4))
    base = rng.uniform(60, 180, size=3).astype(np.float32)
    for c in range(3):
        img[..., c] = base[c]
    for _ in range(ngrad):
        col = rng.uniform(-70, 70, size=3).astype(np.float32)
        if rng.random() < 0.5:
            ang = rng.uniform(0, 2 * np.pi)
            g = np.cos(ang) * xx + np.sin(ang) * yy
        else:
            cx, cy = rng.uniform(0, 1, size=2)
            g = np.sqrt((xx - cx) ** 2 + (yy - cy) ** 2)
        for c in range(3):
            img[..., c] += col[c] * g
    beta = float(rng.uniform(1.6, 2.2))
    amp = float(rng.uniform(18, 40))
    shared = _texture(rng, h, w, beta)
    for c in range(3):
        own = _texture(rng, h, w, beta)
        img[..., c] += amp * (0.9 * shared + 0.436 * own)      # inter-channel correlation ~0.9
    # fine detail: a periodic full-resolution 1/f tile (FFT textures wrap seamlessly)
    fine_amp = float(rng.uniform(6, 26))
    th, tw = min(h, 1024), min(w, 1024)
    tile = _texture(rng, th, tw, float(rng.uniform(0.9, 1.5)))
    fine = np.tile(tile, (-(-h // th), -(-w // tw)))[:h, :w]
    for c in range(3):
        img[..., c] += fine_amp * fine
    sigma = float(rng.uniform(1.5, 6.0))
    img += rng.standard_normal(img.shape, dtype=np.float32) * sigma
    out = np.clip(img + 0.5, 0, 255).astype(np.uint8)
    if channels == 4:
        alpha = np.clip(255.0 * (0.55 + 0.45 * _texture(rng, h, w, 2.0)), 0, 255).astype(np.uint8)
        out = np.dstack([out, alpha])
    return out


def noise_image(index: int, w: int, h: int) -> np.ndarray:
    """Uniform noise: the 0xFF byte-stuffing stress case at high quality."""
    rng = np.random.default_rng(SEED_BASE ^ 0x5A5A ^ index)
    return rng.integers(0, 256, size=(h, w, 3), dtype=np.uint8)


def pattern_image(kind: str, w: int, h: int) -> np.ndarray:
    """Small known-answer patterns: solid, gradient, checker."""
    if kind == "solid":
        out = np.empty((h, w, 3), np.uint8)
        out[...] = (200, 30, 90)
        return out
    if kind == "gradient":
        x = np.arange(w, dtype=np.int32)[None, :]
        y = np.arange(h, dtype=np.int32)[:, None]
        return np.stack([(x * 255 // max(1, w - 1)) + 0 * y, (y * 255 // max(1, h - 1)) + 0 * x,
                         ((x + y) * 255 // max(1, w + h - 2))], axis=-1).astype(np.uint8)
    if kind == "checker":
        x = np.arange(w)[None, :]
        y = np.arange(h)[:, None]
        m = (((x // 3) + (y // 5)) & 1).astype(np.uint8)
        return np.stack([m * 255, 255 - m * 255, m * 128 + 60], axis=-1).astype(np.uint8)
    raise ValueError(kind)
