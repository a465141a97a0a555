"""ctypes loader for the CPU oracle (TEST INFRASTRUCTURE ONLY -- see oracle.h).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
legs may import this module.  The product package must never import it.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_DIR = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_DIR, "liboracle.so")

FMT_RGB8, FMT_BGR8, FMT_RGBA8, FMT_ABGR8, FMT_GRAY8, FMT_ARGB32, FMT_XRGB32 = 0, 1, 2, 3, 4, 5, 6


class Profile(C.Structure):
    _fields_ = [("table_mode", C.c_int), ("jfif_minor", C.c_int)]


JDK = Profile(0, 2)          # float32 table formula, JFIF 1.02  [JDK-mem]
IJG = Profile(1, 1)          # integer jpeg_quality_scaling, JFIF 1.01


class Geom(C.Structure):
    _fields_ = [("w", C.c_int), ("h", C.c_int), ("mcu_cols", C.c_int), ("mcu_rows", C.c_int),
                ("yblk_w", C.c_int), ("yblk_h", C.c_int), ("cblk_w", C.c_int), ("cblk_h", C.c_int),
                ("nblocks", C.c_long)]


def _has_avx2() -> bool:
    try:
        return " avx2 " in open("/proc/cpuinfo").read().replace("\n", " ")
    except OSError:
        return False


def build_flags() -> str:
    """How the oracle is compiled on this host (quoted in bench.py's cpu_baseline.sample)."""
    return "gcc -O3 -march=x86-64-v3, scalar C" if _has_avx2() else "gcc -O3, scalar C"


def build(force: bool = False) -> str:
    srcs = [os.path.join(_DIR, f) for f in ("jpeg_oracle.c", "resize_oracle.c", "jpeg_decode_oracle.c", "oracle.h",
                                            "jpeg_tables.h", "Makefile")]
    stale = force or not os.path.exists(_SO) or any(os.path.getmtime(s) > os.path.getmtime(_SO) for s in srcs)
    if stale:
        cmd = ["make", "-C", _DIR, "-s", "-B", "liboracle.so"]
        if not _has_avx2():
            cmd.append("ARCH=")                    # a host without AVX2: plain x86-64 code
        subprocess.run(cmd, check=True)
    return _SO


_lib = None


def lib():
    global _lib
    if _lib is None:
        _lib = C.CDLL(build())
        u8p, i16p = C.POINTER(C.c_uint8), C.POINTER(C.c_int16)
        L = _lib
        L.orc_build_qtables.argtypes = [C.c_float, C.POINTER(Profile), u8p, u8p]
        L.orc_build_qtables.restype = None
        L.orc_geometry.argtypes = [C.c_int, C.c_int, C.POINTER(Geom)]
        L.orc_geometry.restype = None
        L.orc_forward_coefficients.argtypes = [u8p, C.c_int, C.c_int, C.c_int, C.c_int, i16p]
        L.orc_jpeg_encode_tables.argtypes = [u8p, C.c_int, C.c_int, C.c_int, C.c_int, u8p, u8p, C.c_int,
                                             u8p, C.c_size_t, C.POINTER(C.c_size_t)]
        L.orc_jpeg_encode.argtypes = [u8p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_float, C.POINTER(Profile),
                                      u8p, C.c_size_t, C.POINTER(C.c_size_t)]
        L.orc_find_best_quality.argtypes = [u8p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_long, C.c_float,
                                            C.POINTER(Profile), C.POINTER(C.c_float), C.POINTER(C.c_long),
                                            C.POINTER(C.c_int)]
        L.orc_find_best_quality.restype = C.c_float
        L.orc_jpg_fit.argtypes = [u8p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_float, C.c_long,
                                  C.c_int, C.c_float, C.c_double, C.POINTER(Profile),
                                  u8p, C.c_size_t, C.POINTER(C.c_size_t),
                                  C.POINTER(C.c_float), C.POINTER(C.c_double), C.POINTER(C.c_int),
                                  C.POINTER(C.c_int)]
        L.orc_resize_dims.argtypes = [C.c_int, C.c_int, C.c_double, C.POINTER(C.c_int), C.POINTER(C.c_int)]
        L.orc_resize_dims.restype = None
        for f in (L.orc_resize_bilinear, L.orc_resize_bilinear_f64):
            f.argtypes = [u8p, C.c_int, C.c_int, C.c_int, C.c_int, u8p, C.c_int, C.c_int, C.c_int]
        L.orc_subsample_factor.argtypes = [C.c_int, C.c_int]
        L.orc_subsample.argtypes = [u8p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, u8p, C.c_int]
        L.orc_png_target_dims.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_int),
                                          C.POINTER(C.c_int), C.POINTER(C.c_double)]
        L.orc_jpeg_info.argtypes = [u8p, C.c_size_t, C.POINTER(C.c_int), C.POINTER(C.c_int)]
        L.orc_jpeg_decode_coefficients.argtypes = [u8p, C.c_size_t, i16p]
        L.orc_jpeg_decode.argtypes = [u8p, C.c_size_t, C.c_int, u8p, C.c_int]
        L.orc_jpeg_sampling.argtypes = [u8p, C.c_size_t, C.POINTER(C.c_int), C.POINTER(C.c_int)]
    return _lib


def _u8(a: np.ndarray):
    return a.ctypes.data_as(C.POINTER(C.c_uint8))


def _img(img: np.ndarray):
    img = np.ascontiguousarray(img, dtype=np.uint8)
    if img.ndim == 2:
        img = img[:, :, None]
    assert img.ndim == 3 and img.shape[2] in (1, 3, 4)
    h, w, c = img.shape
    return img, w, h, w * c


def qtables(q: float, prof: Profile = JDK):
    lum = np.zeros(64, np.uint8)
    chr_ = np.zeros(64, np.uint8)
    lib().orc_build_qtables(C.c_float(q), C.byref(prof), _u8(lum), _u8(chr_))
    return lum, chr_


def geometry(w: int, h: int) -> Geom:
    g = Geom()
    lib().orc_geometry(w, h, C.byref(g))
    return g


def forward_coefficients(img: np.ndarray, fmt: int = FMT_RGB8) -> np.ndarray:
    img, w, h, pitch = _img(img)
    g = geometry(w, h)
    out = np.zeros((g.nblocks, 64), np.int16)
    rc = lib().orc_forward_coefficients(_u8(img), w, h, pitch, fmt, out.ctypes.data_as(C.POINTER(C.c_int16)))
    assert rc == 0, rc
    return out


def encode_tables(img: np.ndarray, lum: np.ndarray, chr_: np.ndarray, jfif_minor: int = 2,
                  fmt: int = FMT_RGB8) -> bytes:
    img, w, h, pitch = _img(img)
    lum = np.ascontiguousarray(lum, np.uint8)
    chr_ = np.ascontiguousarray(chr_, np.uint8)
    n = C.c_size_t(0)
    cap = w * h * 3 + 65536
    out = np.zeros(cap, np.uint8)
    rc = lib().orc_jpeg_encode_tables(_u8(img), w, h, pitch, fmt, _u8(lum), _u8(chr_), jfif_minor,
                                      _u8(out), cap, C.byref(n))
    assert rc == 0 and n.value <= cap, (rc, n.value)
    return out[: n.value].tobytes()


def encode(img: np.ndarray, q: float, prof: Profile = JDK, fmt: int = FMT_RGB8) -> bytes:
    lum, chr_ = qtables(q, prof)
    return encode_tables(img, lum, chr_, prof.jfif_minor, fmt)


def encoded_size(img: np.ndarray, q: float, prof: Profile = JDK, fmt: int = FMT_RGB8) -> int:
    img, w, h, pitch = _img(img)
    n = C.c_size_t(0)
    rc = lib().orc_jpeg_encode(_u8(img), w, h, pitch, fmt, C.c_float(q), C.byref(prof), None, 0, C.byref(n))
    assert rc == 0, rc
    return n.value


def find_best_quality(img: np.ndarray, target: int, initial_q: float, prof: Profile = JDK, fmt: int = FMT_RGB8):
    img, w, h, pitch = _img(img)
    tq = (C.c_float * 8)()
    ts = (C.c_long * 8)()
    nt = C.c_int(0)
    best = lib().orc_find_best_quality(_u8(img), w, h, pitch, fmt, target, C.c_float(initial_q), C.byref(prof),
                                       tq, ts, C.byref(nt))
    return float(best), [(float(tq[i]), int(ts[i])) for i in range(nt.value)]


def jpg_fit(img: np.ndarray, initial_q: float, target: int, cached=None, prof: Profile = JDK,
            fmt: int = FMT_RGB8):
    """-> (status, bytes, best_q, scale, used_cache, n_encodes); status 0 fits / 1 no fit."""
    img, w, h, pitch = _img(img)
    cap = max(int(target), 0) + 4096
    out = np.zeros(cap, np.uint8)
    n = C.c_size_t(0)
    bq = C.c_float(0)
    sc = C.c_double(0)
    uc = C.c_int(0)
    ne = C.c_int(0)
    has = 1 if cached is not None else 0
    cq, cs = cached if cached is not None else (0.0, 1.0)
    rc = lib().orc_jpg_fit(_u8(img), w, h, pitch, fmt, C.c_float(initial_q), int(target), has, C.c_float(cq),
                           C.c_double(cs), C.byref(prof), _u8(out), cap, C.byref(n), C.byref(bq), C.byref(sc),
                           C.byref(uc), C.byref(ne))
    return rc, out[: min(n.value, cap)].tobytes(), float(bq.value), float(sc.value), int(uc.value), int(ne.value)


def resize_dims(w: int, h: int, scale: float):
    a, b = C.c_int(0), C.c_int(0)
    lib().orc_resize_dims(w, h, C.c_double(scale), C.byref(a), C.byref(b))
    return a.value, b.value


def png_target_dims(w: int, h: int, min_w: int, min_h: int):
    a, b, s = C.c_int(0), C.c_int(0), C.c_double(0)
    rc = lib().orc_png_target_dims(w, h, min_w, min_h, C.byref(a), C.byref(b), C.byref(s))
    return rc, a.value, b.value, s.value


def subsample_factor(w: int, h: int) -> int:
    return lib().orc_subsample_factor(w, h)


def subsample(img: np.ndarray, s: int) -> np.ndarray:
    img, w, h, pitch = _img(img)
    c = img.shape[2]
    dst = np.zeros(((h + s - 1) // s, (w + s - 1) // s, c), np.uint8)
    rc = lib().orc_subsample(_u8(img), w, h, pitch, c, s, _u8(dst), dst.strides[0])
    assert rc == 0, rc
    return dst


def resize(img: np.ndarray, dw: int, dh: int, fmt: int = FMT_RGB8, f64: bool = False) -> np.ndarray:
    img, w, h, pitch = _img(img)
    c = img.shape[2]
    dst = np.zeros((dh, dw, c), np.uint8)
    fn = lib().orc_resize_bilinear_f64 if f64 else lib().orc_resize_bilinear
    rc = fn(_u8(img), w, h, pitch, fmt, _u8(dst), dw, dh, dw * c)
    assert rc == 0, rc
    return dst


# ---- f2: the JPEG reader of core/ImageCompression.java:119-155 (jpeg_decode_oracle.c) ----
DEC_UNSUPPORTED = -4


def jpeg_info(data: bytes):
    """(w, h) of a file the decoder takes; raises ValueError(code) for unsupported / corrupt files."""
    buf = np.frombuffer(data, np.uint8)
    w, h = C.c_int(0), C.c_int(0)
    rc = lib().orc_jpeg_info(_u8(buf), buf.size, C.byref(w), C.byref(h))
    if rc:
        raise ValueError(rc)
    return w.value, h.value


def jpeg_decode(data: bytes, fmt: int = FMT_RGB8) -> np.ndarray:
    w, h = jpeg_info(data)
    buf = np.frombuffer(data, np.uint8)
    out = np.empty((h, w, 3), np.uint8)
    rc = lib().orc_jpeg_decode(_u8(buf), buf.size, fmt, _u8(out), w * 3)
    if rc:
        raise ValueError(rc)
    return out


def jpeg_decode_coefficients(data: bytes) -> np.ndarray:
    """Huffman-decoded coefficients, [6 * mcus][64] int16, natural order, DC prediction undone, not dequantised."""
    w, h = jpeg_info(data)
    buf = np.frombuffer(data, np.uint8)
    hs, vs = C.c_int(0), C.c_int(0)
    lib().orc_jpeg_sampling(_u8(buf), buf.size, C.byref(hs), C.byref(vs))
    nb = (hs.value * vs.value + 2) * ((w + 8 * hs.value - 1) // (8 * hs.value)) * ((h + 8 * vs.value - 1) // (8 * vs.value))
    out = np.empty((nb, 64), np.int16)
    rc = lib().orc_jpeg_decode_coefficients(_u8(buf), buf.size, out.ctypes.data_as(C.POINTER(C.c_int16)))
    if rc:
        raise ValueError(rc)
    return out
