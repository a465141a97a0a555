"""ctypes binding of libimgcomp_b200.so plus a host-side mirror of the reference's
four seams (same names, argument meaning and error behaviour as the Java statics;
see package docstring).  The Java 21 FFM binding a maintainer would add is in
INTEGRATION.md; this module is the same binding for the pytest harness / bench.
"""
from __future__ import annotations

import ctypes as C
import os
from dataclasses import dataclass
from typing import Dict, MutableMapping, Optional, Tuple

import numpy as np

ICB_OK, ICB_NO_FIT = 0, 1
ICB_ERR_INVALID, ICB_ERR_OOM, ICB_ERR_CUDA, ICB_ERR_UNSUPPORTED = -1, -2, -3, -4
ICB_ERR_CAPACITY, ICB_ERR_NO_DEVICE, ICB_ERR_TICKET = -5, -6, -7
FMT_RGB8, FMT_BGR8, FMT_RGBA8, FMT_ABGR8, FMT_GRAY8, FMT_ARGB32, FMT_XRGB32 = 0, 1, 2, 3, 4, 5, 6
_BPP = {FMT_RGB8: 3, FMT_BGR8: 3, FMT_RGBA8: 4, FMT_ABGR8: 4, FMT_GRAY8: 1, FMT_ARGB32: 4, FMT_XRGB32: 4}

_HERE = os.path.dirname(os.path.abspath(__file__))
# IMGCOMP_B200_LIB: load another build of the SAME library (A/B runs of kernel variants); never a fallback
_LIB_PATH = os.environ.get("IMGCOMP_B200_LIB") or os.path.join(_HERE, "libimgcomp_b200.so")

# every symbol include/imgcomp_b200.h declares (checked by tests/test_abi.py)
ABI_SYMBOLS = [
    "icb_init", "icb_shutdown", "icb_device_count", "icb_last_error", "icb_set_profile",
    "icb_alloc_pinned", "icb_free_pinned", "icb_jpg_fit", "icb_jpg_encode", "icb_jpg_sizes",
    "icb_jpg_coefficients", "icb_resize_dims", "icb_resize_bilinear", "icb_png_fit", "icb_png_target_dims",
    "icb_subsample_factor", "icb_subsample",
    "icb_submit_jpg_fit", "icb_wait", "icb_poll", "icb_submit_resize", "icb_submit_png_fit",
    "icb_png_encode", "icb_submit_png_encode", "icb_png_compress", "icb_submit_png_compress",
    "icb_profile_resize_series", "icb_png_encode_device_noalloc",
    "icb_jpg_file_info", "icb_jpg_decode", "icb_jpg_file_fit", "icb_submit_jpg_file_fit",
    "icb_run_list",
    "icb_image_upload", "icb_image_free",
    "icb_submit_jpg_fit_image", "icb_submit_resize_image", "icb_stats", "icb_profile_jpg_fit_image",
    "icb_profile_transform_series",
    "icb_mark", "icb_elapsed_ms",
]


class IcbError(RuntimeError):
    def __init__(self, code: int, msg: str = ""):
        super().__init__(f"imgcomp_b200 status {code}: {msg}")
        self.code = code


class _Result(C.Structure):
    _fields_ = [("status", C.c_int), ("out_len", C.c_size_t), ("best_q", C.c_float), ("scale", C.c_double),
                ("used_cache", C.c_int), ("n_passes", C.c_int), ("device", C.c_int)]


class _Stats(C.Structure):
    _fields_ = [("images_done", C.c_uint64), ("kernels_launched", C.c_uint64), ("entropy_passes", C.c_uint64),
                ("h2d_bytes", C.c_uint64), ("d2h_bytes", C.c_uint64), ("last_transform_ms", C.c_float),
                ("last_entropy_ms", C.c_float), ("last_total_ms", C.c_float),
                ("last_transform_launches", C.c_int), ("last_entropy_launches", C.c_int),
                ("kernel_clock_ns", C.c_uint64 * 3), ("kernel_clock_launches", C.c_uint64 * 3)]


class _Item(C.Structure):
    """icb_item (include/imgcomp_b200.h)"""
    _fields_ = [("kind", C.c_int), ("px", C.c_void_p), ("img", C.c_void_p), ("w", C.c_int), ("h", C.c_int),
                ("pitch", C.c_int), ("fmt", C.c_int), ("init_q", C.c_float), ("target", C.c_int64),
                ("has_cached", C.c_int), ("cached_q", C.c_float), ("cached_scale", C.c_double), ("min_w", C.c_int),
                ("min_h", C.c_int), ("out", C.c_void_p), ("out_cap", C.c_size_t), ("dw", C.c_int), ("dh", C.c_int),
                ("res", _Result)]


ITEM_KINDS = {"jpg_fit": 0, "jpg_fit_image": 1, "png_fit": 2, "png_compress": 3, "resize_image": 4, "jpg_file_fit": 5}


class ItemList:
    """A prepared array of icb_item for Engine.run_list (built once, run many times: nothing of Python sits
    between the items of a run).  Keeps the numpy buffers it points into alive."""

    def __init__(self, arr, keep):
        self.arr, self._keep = arr, keep

    def __len__(self):
        return len(self.arr)

    def results(self):
        return [FitResult(int(it.res.status), None, float(it.res.best_q), float(it.res.scale), bool(it.res.used_cache),
                          int(it.res.n_passes), int(it.res.device), int(it.res.out_len)) for it in self.arr]

    def dims(self):
        return [(int(it.dw), int(it.dh)) for it in self.arr]


def jpg_file_info(data):
    """(w, h) of a JPEG file the GPU reader takes (host parse only); IcbError otherwise."""
    buf = data if isinstance(data, np.ndarray) else np.frombuffer(data, np.uint8)
    w, h = C.c_int(0), C.c_int(0)
    rc = load_library().icb_jpg_file_info(buf.ctypes.data, buf.size, C.byref(w), C.byref(h))
    if rc:
        raise IcbError(rc, "not a file the GPU JPEG reader takes")
    return w.value, h.value


def library_path() -> str:
    return _LIB_PATH


_lib = None


def load_library(build_if_missing: bool = True):
    """dlopen the CUDA library (building it in-tree with nvcc if it is absent).
    Raises if it cannot be had -- there is no other implementation to fall back to."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(_LIB_PATH):
        if not build_if_missing:
            raise IcbError(ICB_ERR_NO_DEVICE, f"{_LIB_PATH} not built")
        from . import build as _build
        _build.build()
    L = C.CDLL(_LIB_PATH)
    u8p, vp = C.POINTER(C.c_uint8), C.c_void_p
    L.icb_init.argtypes = [C.POINTER(C.c_int), C.c_int, C.c_int, C.c_int, C.c_int, C.POINTER(vp)]
    L.icb_shutdown.argtypes = [vp]
    L.icb_shutdown.restype = None
    L.icb_device_count.argtypes = [vp]
    L.icb_last_error.argtypes = [vp]
    L.icb_last_error.restype = C.c_char_p
    L.icb_set_profile.argtypes = [vp, C.c_int, C.c_int]
    L.icb_alloc_pinned.argtypes = [C.c_size_t]
    L.icb_alloc_pinned.restype = vp
    L.icb_free_pinned.argtypes = [vp]
    L.icb_free_pinned.restype = None
    L.icb_jpg_fit.argtypes = [vp, vp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_float, C.c_int64, C.c_int, C.c_float,
                              C.c_double, vp, C.c_size_t, C.POINTER(C.c_size_t), C.POINTER(C.c_float),
                              C.POINTER(C.c_double), C.POINTER(C.c_int)]
    L.icb_jpg_encode.argtypes = [vp, vp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_float, vp, C.c_size_t,
                                 C.POINTER(C.c_size_t)]
    L.icb_jpg_sizes.argtypes = [vp, vp, C.c_int, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_float), C.c_int,
                                C.POINTER(C.c_int64)]
    L.icb_jpg_coefficients.argtypes = [vp, vp, C.c_int, C.c_int, C.c_int, C.c_int, vp, C.c_size_t]
    L.icb_resize_dims.argtypes = [C.c_int, C.c_int, C.c_double, C.POINTER(C.c_int), C.POINTER(C.c_int)]
    L.icb_resize_dims.restype = None
    L.icb_resize_bilinear.argtypes = [vp, vp, C.c_int, C.c_int, C.c_int, C.c_int, vp, C.c_int, C.c_int, C.c_int]
    L.icb_png_fit.argtypes = [vp, vp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, vp, C.c_size_t,
                              C.POINTER(C.c_int), C.POINTER(C.c_int)]
    L.icb_png_target_dims.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_int),
                                      C.POINTER(C.c_double)]
    L.icb_subsample_factor.argtypes = [C.c_int, C.c_int]
    L.icb_subsample.argtypes = [vp, vp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, vp, C.c_int]
    L.icb_submit_jpg_fit.argtypes = [vp, vp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_float, C.c_int64, C.c_int,
                                     C.c_float, C.c_double, vp, C.c_size_t, C.POINTER(C.c_int64)]
    L.icb_wait.argtypes = [vp, C.c_int64, C.POINTER(_Result)]
    L.icb_poll.argtypes = [vp, C.c_int64, C.POINTER(C.c_int)]
    L.icb_submit_resize.argtypes = [vp, vp, C.c_int, C.c_int, C.c_int, C.c_int, vp, C.c_int, C.c_int, C.c_int,
                                    C.POINTER(C.c_int64)]
    L.icb_submit_png_fit.argtypes = [vp, vp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, vp, C.c_size_t,
                                     C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_int64)]
    L.icb_png_encode.argtypes = [vp, vp, C.c_int, C.c_int, C.c_int, C.c_int, vp, C.c_size_t, C.POINTER(C.c_size_t)]
    L.icb_submit_png_encode.argtypes = [vp, vp, C.c_int, C.c_int, C.c_int, C.c_int, vp, C.c_size_t,
                                        C.POINTER(C.c_int64)]
    L.icb_png_compress.argtypes = [vp, vp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, vp, C.c_size_t,
                                   C.POINTER(C.c_size_t), C.POINTER(C.c_int), C.POINTER(C.c_int)]
    L.icb_submit_png_compress.argtypes = [vp, vp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, vp,
                                          C.c_size_t, C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_int64)]
    L.icb_png_encode_device_noalloc.argtypes = [vp, vp, C.c_int, C.c_int, C.c_int, C.c_int, vp, C.c_size_t, vp, vp]
    L.icb_jpg_file_info.argtypes = [vp, C.c_size_t, C.POINTER(C.c_int), C.POINTER(C.c_int)]
    L.icb_jpg_decode.argtypes = [vp, vp, C.c_size_t, C.c_int, vp, C.c_size_t, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_int)]
    L.icb_jpg_file_fit.argtypes = [vp, vp, C.c_size_t, C.c_float, C.c_int64, C.c_int, C.c_float, C.c_double, vp, C.c_size_t,
                                   C.POINTER(C.c_size_t), C.POINTER(C.c_float), C.POINTER(C.c_double), C.POINTER(C.c_int),
                                   C.POINTER(C.c_int), C.POINTER(C.c_int)]
    L.icb_submit_jpg_file_fit.argtypes = [vp, vp, C.c_size_t, C.c_float, C.c_int64, C.c_int, C.c_float, C.c_double, vp,
                                          C.c_size_t, C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_int64)]
    L.icb_profile_resize_series.argtypes = [vp, vp, C.c_int, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_float)]
    L.icb_run_list.argtypes = [vp, C.POINTER(_Item), C.c_size_t, C.c_int, C.c_int]
    L.icb_image_upload.argtypes = [vp, C.c_int, vp, C.c_int, C.c_int, C.c_int, C.c_int, C.POINTER(vp)]
    L.icb_image_free.argtypes = [vp, vp]
    L.icb_image_free.restype = None
    L.icb_submit_jpg_fit_image.argtypes = [vp, vp, C.c_float, C.c_int64, C.c_int, C.c_float, C.c_double, vp,
                                           C.c_size_t, C.POINTER(C.c_int64)]
    L.icb_submit_resize_image.argtypes = [vp, vp, C.c_int, C.c_int, vp, C.c_int, C.POINTER(C.c_int64)]
    L.icb_stats.argtypes = [vp, C.POINTER(_Stats)]
    L.icb_profile_jpg_fit_image.argtypes = [vp, vp, C.c_float, C.c_int64, C.c_int, C.POINTER(C.c_float),
                                            C.POINTER(C.c_float), C.POINTER(C.c_float), C.POINTER(C.c_float),
                                            C.POINTER(C.c_int)]
    L.icb_profile_transform_series.argtypes = [vp, C.POINTER(vp), C.c_int, C.c_int, C.c_float, C.POINTER(C.c_float),
                                               C.POINTER(C.c_int)]
    L.icb_mark.argtypes = [vp, C.c_int, C.c_int]
    L.icb_elapsed_ms.argtypes = [vp, C.c_int, C.POINTER(C.c_float)]
    _lib = L
    return L


# ---- value types at the seam (report/CompressionParams.java:3, learn/LearnedParams.java:8,
#      learn/jpg/SimilarityKey.java:9) ----
@dataclass(frozen=True)
class CompressionParams:
    quality: float
    min_size_bytes: int
    min_width: int
    min_height: int
    target_max_size_bytes: int


@dataclass(frozen=True)
class LearnedParams:
    quality: float
    scale: float


@dataclass(frozen=True)
class SimilarityKey:
    width_bucket: int
    height_bucket: int
    size_bucket: int


def create_key(width: int, height: int, file_size: int) -> SimilarityKey:
    """tools/CacheTools.java:14-21 -- buckets of 100 px / 100 px / 100 KiB on the DECODED dims."""
    return SimilarityKey(width // 100, height // 100, file_size // 102400)


@dataclass
class FitResult:
    status: int
    data: Optional[bytes]
    best_q: float
    scale: float
    used_cache: bool
    n_passes: int
    device: int
    out_len: int = 0


def _image_args(image: np.ndarray, fmt: Optional[int]) -> Tuple[np.ndarray, int, int, int, int]:
    if image is None:
        raise TypeError("image must not be None")         # Java: NullPointerException
    if image.dtype == np.uint8 and image.ndim == 2:
        image = image[:, :, None]                         # TYPE_BYTE_GRAY
    if image.dtype != np.uint8 or image.ndim != 3 or image.shape[2] not in (1, 3, 4):
        raise ValueError("image must be uint8 (h, w[, 1|3|4])")
    if not image.flags["C_CONTIGUOUS"]:
        image = np.ascontiguousarray(image)
    h, w, c = image.shape
    if fmt is None:
        fmt = {1: FMT_GRAY8, 3: FMT_RGB8, 4: FMT_RGBA8}[c]
    if _BPP.get(fmt) != c:
        raise ValueError("pixel format does not match the channel count")
    return image, w, h, image.strides[0], fmt


def resize_dims(w: int, h: int, scale: float) -> Tuple[int, int]:
    """tools/ImageTools.java:8-9 (host arithmetic in the library, no device needed)."""
    L = load_library()
    a, b = C.c_int(0), C.c_int(0)
    L.icb_resize_dims(w, h, C.c_double(scale), C.byref(a), C.byref(b))
    return a.value, b.value


def subsample_factor(w: int, h: int) -> int:
    """core/ImageCompression.java:140-153: the integer source subsampling the reference decodes
    with (1 when the longer side is at most 4096); host arithmetic, no device needed."""
    return load_library().icb_subsample_factor(w, h)


def png_target_dims(w: int, h: int, min_w: int, min_h: int):
    """core/ImageCompressionPng.java:49-61 -> None when no resize is due (the seam returns false),
    else (new_w, new_h, scale)."""
    L = load_library()
    a, b, s = C.c_int(0), C.c_int(0), C.c_double(0)
    rc = L.icb_png_target_dims(w, h, min_w, min_h, C.byref(a), C.byref(b), C.byref(s))
    if rc == ICB_NO_FIT:
        return None
    if rc:
        raise IcbError(rc, "bad dimensions")
    return a.value, b.value, s.value


class Engine:
    """One handle = the GPU work queues that replace CompressionBatch's thread pool
    (CompressionBatch.java:64-68).  Thread-safe: call from as many threads as slots."""

    def __init__(self, devices=None, slots_per_gpu: int = 4, max_w: int = 0, max_h: int = 0):
        self._L = load_library()
        self._h = C.c_void_p()
        arr, n = None, 0
        if devices:
            arr = (C.c_int * len(devices))(*devices)
            n = len(devices)
        rc = self._L.icb_init(arr, n, slots_per_gpu, max_w, max_h, C.byref(self._h))
        if rc != ICB_OK:
            self._h = C.c_void_p()
            raise IcbError(rc, "icb_init failed (no CUDA device? this library has no CPU path)")
        self._pinned = {}
        self._inflight = {}

    # ---- lifecycle ----
    def close(self):
        if getattr(self, "_h", None) and self._h.value:
            for p in list(self._pinned.values()):
                self._L.icb_free_pinned(p)
            self._pinned.clear()
            self._L.icb_shutdown(self._h)
            self._h = C.c_void_p()

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _err(self, rc: int) -> IcbError:
        return IcbError(rc, (self._L.icb_last_error(self._h) or b"").decode(errors="replace"))

    @property
    def device_count(self) -> int:
        return self._L.icb_device_count(self._h)

    def set_profile(self, table_mode: int = 0, jfif_minor: int = 2):
        rc = self._L.icb_set_profile(self._h, table_mode, jfif_minor)
        if rc:
            raise self._err(rc)

    def alloc_pinned(self, shape, dtype=np.uint8) -> np.ndarray:
        nbytes = int(np.prod(shape)) * np.dtype(dtype).itemsize
        p = self._L.icb_alloc_pinned(nbytes)
        if not p:
            raise IcbError(ICB_ERR_OOM, "pinned allocation failed")
        buf = (C.c_uint8 * nbytes).from_address(p)
        arr = np.frombuffer(buf, dtype=dtype).reshape(shape)
        self._pinned[arr.ctypes.data] = p
        return arr

    def free_pinned(self, arr: np.ndarray) -> None:
        p = self._pinned.pop(arr.ctypes.data, None)
        if p:
            self._L.icb_free_pinned(p)

    # ---- core/ImageCompressionJpg.java:136-147 ----
    def compress_jpg_to_stream(self, image: np.ndarray, quality: float, fmt: Optional[int] = None) -> bytes:
        image, w, h, pitch, fmt = _image_args(image, fmt)
        cap = w * h * 3 + 65536
        while True:
            out = np.empty(cap, np.uint8)
            n = C.c_size_t(0)
            rc = self._L.icb_jpg_encode(self._h, image.ctypes.data, w, h, pitch, fmt, C.c_float(quality),
                                        out.ctypes.data, cap, C.byref(n))
            if rc == ICB_ERR_CAPACITY and n.value > cap:
                cap = n.value
                continue
            if rc:
                raise self._err(rc)
            return out[: n.value].tobytes()

    # ---- core/ImageCompressionJpg.java:77-122 (+ :216-238, :158-200, :255-260) ----
    def jpg_fit(self, image: np.ndarray, init_q: float, target: int, cached: Optional[LearnedParams] = None,
                fmt: Optional[int] = None) -> FitResult:
        image, w, h, pitch, fmt = _image_args(image, fmt)
        cap = max(int(target), 0) + 16
        out = np.empty(cap, np.uint8)
        n, bq, sc, uc = C.c_size_t(0), C.c_float(0), C.c_double(0), C.c_int(0)
        rc = self._L.icb_jpg_fit(self._h, image.ctypes.data, w, h, pitch, fmt, C.c_float(init_q), int(target),
                                 1 if cached else 0, C.c_float(cached.quality if cached else 0.0),
                                 C.c_double(cached.scale if cached else 1.0), out.ctypes.data, cap, C.byref(n),
                                 C.byref(bq), C.byref(sc), C.byref(uc))
        if rc < 0:
            raise self._err(rc)
        data = out[: n.value].tobytes() if rc == ICB_OK else None
        return FitResult(rc, data, float(bq.value), float(sc.value), bool(uc.value), 0, 0, n.value)

    def compress_jpg_with_target_size(self, original_image: np.ndarray, original_size: int, output_file,
                                      params: CompressionParams,
                                      cache: MutableMapping[SimilarityKey, LearnedParams],
                                      fmt: Optional[int] = None) -> bool:
        """Same contract as the Java static: returns True and writes output_file when a
        (quality, scale) fits params.target_max_size_bytes, updating `cache` on a fresh search."""
        if original_image is None or params is None or cache is None:
            raise TypeError("null argument")
        h, w = original_image.shape[:2]
        key = create_key(w, h, original_size)                                   # :79
        cached = cache.get(key)                                                 # :80
        r = self.jpg_fit(original_image, params.quality, params.target_max_size_bytes, cached, fmt)
        if r.status != ICB_OK:
            return False                                                        # :120-121
        if output_file is not None:
            with open(output_file, "wb") as f:                                  # :258 / :229
                f.write(r.data)
        if not r.used_cache:
            cache[key] = LearnedParams(r.best_q, r.scale)                       # :111
        self.last_fit = r
        return True

    # ---- hooks ----
    def jpg_sizes(self, image: np.ndarray, qualities, fmt: Optional[int] = None):
        image, w, h, pitch, fmt = _image_args(image, fmt)
        qs = np.ascontiguousarray(qualities, np.float32)
        sizes = np.zeros(len(qs), np.int64)
        rc = self._L.icb_jpg_sizes(self._h, image.ctypes.data, w, h, pitch, fmt,
                                   qs.ctypes.data_as(C.POINTER(C.c_float)), len(qs),
                                   sizes.ctypes.data_as(C.POINTER(C.c_int64)))
        if rc:
            raise self._err(rc)
        return sizes

    def jpg_coefficients(self, image: np.ndarray, fmt: Optional[int] = None) -> np.ndarray:
        image, w, h, pitch, fmt = _image_args(image, fmt)
        nblocks = 6 * ((w + 15) // 16) * ((h + 15) // 16)
        out = np.zeros((nblocks, 64), np.int16)
        rc = self._L.icb_jpg_coefficients(self._h, image.ctypes.data, w, h, pitch, fmt, out.ctypes.data, out.size)
        if rc:
            raise self._err(rc)
        return out

    # ---- tools/ImageTools.java:7-26 ----
    def resize_dims(self, w: int, h: int, scale: float) -> Tuple[int, int]:
        a, b = C.c_int(0), C.c_int(0)
        self._L.icb_resize_dims(w, h, C.c_double(scale), C.byref(a), C.byref(b))
        return a.value, b.value

    def resize_image(self, original_image: np.ndarray, scale: float, fmt: Optional[int] = None) -> np.ndarray:
        image, w, h, pitch, fmt = _image_args(original_image, fmt)
        dw, dh = self.resize_dims(w, h, scale)
        return self.resize_to(image, dw, dh, fmt)

    def resize_to(self, image: np.ndarray, dw: int, dh: int, fmt: Optional[int] = None) -> np.ndarray:
        image, w, h, pitch, fmt = _image_args(image, fmt)
        c = image.shape[2]
        dst = np.empty((dh, dw, c), np.uint8)
        rc = self._L.icb_resize_bilinear(self._h, image.ctypes.data, w, h, pitch, fmt, dst.ctypes.data, dw, dh,
                                         dw * c)
        if rc:
            raise self._err(rc)
        return dst

    def resize_to_into(self, image: np.ndarray, dst: np.ndarray, fmt: Optional[int] = None) -> np.ndarray:
        """Same, into a caller-owned (e.g. pinned) destination of shape (dh, dw, c)."""
        image, w, h, pitch, fmt = _image_args(image, fmt)
        if dst.dtype != np.uint8 or dst.ndim != 3 or dst.shape[2] != image.shape[2]:
            raise ValueError("dst must be uint8 (dh, dw, c)")
        rc = self._L.icb_resize_bilinear(self._h, image.ctypes.data, w, h, pitch, fmt, dst.ctypes.data,
                                         dst.shape[1], dst.shape[0], dst.strides[0])
        if rc:
            raise self._err(rc)
        return dst

    # ---- input staging: core/ImageCompression.java:148-153 setSourceSubsampling(s, s, 0, 0) ----
    def subsample(self, image: np.ndarray, s: int, fmt: Optional[int] = None) -> np.ndarray:
        image, w, h, pitch, fmt = _image_args(image, fmt)
        c = image.shape[2]
        dst = np.empty(((h + s - 1) // s, (w + s - 1) // s, c), np.uint8)
        rc = self._L.icb_subsample(self._h, image.ctypes.data, w, h, pitch, fmt, s, dst.ctypes.data, dst.strides[0])
        if rc:
            raise self._err(rc)
        return dst

    # ---- core/ImageCompressionPng.java:37-75 (PNG file write stays with the caller) ----
    def compress_png_with_target_size(self, original_image: np.ndarray, output_file, params: CompressionParams,
                                      fmt: Optional[int] = None):
        """-> (False, None) when the image already fits min_width x min_height (:49-53),
        else (True, resized pixels).  None arguments raise TypeError (Java: NPE, :39-41)."""
        if original_image is None:
            raise TypeError("originalImage must not be null")
        if output_file is None:
            raise TypeError("outputFile must not be null")
        if params is None:
            raise TypeError("params must not be null")
        image, w, h, pitch, fmt = _image_args(original_image, fmt)
        c = image.shape[2]
        dw, dh = C.c_int(0), C.c_int(0)
        rc = self._L.icb_png_fit(self._h, image.ctypes.data, w, h, pitch, fmt, params.min_width, params.min_height,
                                 None, 0, C.byref(dw), C.byref(dh))
        if rc == ICB_NO_FIT:
            return False, None
        if rc:
            raise self._err(rc)
        dst = np.empty((dh.value, dw.value, c), np.uint8)
        rc = self._L.icb_png_fit(self._h, image.ctypes.data, w, h, pitch, fmt, params.min_width, params.min_height,
                                 dst.ctypes.data, dst.size, C.byref(dw), C.byref(dh))
        if rc:
            raise self._err(rc)
        return True, dst

    # ---- core/ImageCompressionPng.java:70, and the whole seam :37-75, with the PNG file made on the GPU ----
    def png_encode(self, image: np.ndarray, fmt: Optional[int] = None) -> bytes:
        """ImageIO.write(image, "png", ...) on the GPU -> the PNG file bytes (decode to exactly `image`)."""
        image, w, h, pitch, fmt = _image_args(image, fmt)
        cap = image.size + image.size // 4 + 65536
        while True:
            out = np.empty(cap, np.uint8)
            n = C.c_size_t(0)
            rc = self._L.icb_png_encode(self._h, image.ctypes.data, w, h, pitch, fmt, out.ctypes.data, cap, C.byref(n))
            if rc == ICB_ERR_CAPACITY and n.value > cap:
                cap = n.value
                continue
            if rc:
                raise self._err(rc)
            return out[: n.value].tobytes()

    def png_compress(self, image: np.ndarray, min_w: int, min_h: int, fmt: Optional[int] = None):
        """compressPngWithTargetSize as one call: dimension rule, resize, PNG file; the resized pixels never
        leave the GPU.  -> None when the image is already inside the box, else (png bytes, dw, dh)."""
        image, w, h, pitch, fmt = _image_args(image, fmt)
        dims = png_target_dims(w, h, min_w, min_h)
        if dims is None:
            return None
        c = image.shape[2]
        cap = dims[0] * dims[1] * c * 5 // 4 + 65536
        while True:
            out = np.empty(cap, np.uint8)
            n, dw, dh = C.c_size_t(0), C.c_int(0), C.c_int(0)
            rc = self._L.icb_png_compress(self._h, image.ctypes.data, w, h, pitch, fmt, min_w, min_h, out.ctypes.data,
                                          cap, C.byref(n), C.byref(dw), C.byref(dh))
            if rc == ICB_ERR_CAPACITY and n.value > cap:
                cap = n.value
                continue
            if rc:
                raise self._err(rc)
            return out[: n.value].tobytes(), dw.value, dh.value

    def submit_png_compress(self, image: np.ndarray, min_w: int, min_h: int, out: np.ndarray, fmt: Optional[int] = None):
        """Queue form -> (ticket, dw, dh) or None; wait(ticket).out_len = bytes of the PNG file in `out`."""
        image, w, h, pitch, fmt = _image_args(image, fmt)
        t, dw, dh = C.c_int64(0), C.c_int(0), C.c_int(0)
        rc = self._L.icb_submit_png_compress(self._h, image.ctypes.data, w, h, pitch, fmt, min_w, min_h,
                                             out.ctypes.data, out.size, C.byref(dw), C.byref(dh), C.byref(t))
        if rc == ICB_NO_FIT:
            return None
        if rc:
            raise self._err(rc)
        self._inflight[t.value] = (image, out)
        return t.value, dw.value, dh.value

    def profile_resize_series(self, img_handle, dw: int, dh: int, reps: int = 20, with_png: bool = False) -> float:
        """ms per launch of the resize kernel (or resize + PNG encoder) on a resident image, back to back."""
        ms = C.c_float(0)
        rc = self._L.icb_profile_resize_series(self._h, img_handle, dw, dh, reps, 1 if with_png else 0, C.byref(ms))
        if rc:
            raise self._err(rc)
        return float(ms.value)

    # ---- asynchronous queue (CompressionBatch.java:77-88 submit / :97-107 await) ----
    def submit_jpg_fit(self, image: np.ndarray, init_q: float, target: int, out: np.ndarray,
                       cached: Optional[LearnedParams] = None, fmt: Optional[int] = None) -> int:
        image, w, h, pitch, fmt = _image_args(image, fmt)
        t = C.c_int64(0)
        rc = self._L.icb_submit_jpg_fit(self._h, image.ctypes.data, w, h, pitch, fmt, C.c_float(init_q), int(target),
                                        1 if cached else 0, C.c_float(cached.quality if cached else 0.0),
                                        C.c_double(cached.scale if cached else 1.0), out.ctypes.data, out.size,
                                        C.byref(t))
        if rc:
            raise self._err(rc)
        self._inflight[t.value] = (image, out)
        return t.value

    def submit_jpg_fit_image(self, img_handle, init_q: float, target: int, out: np.ndarray,
                             cached: Optional[LearnedParams] = None) -> int:
        t = C.c_int64(0)
        rc = self._L.icb_submit_jpg_fit_image(self._h, img_handle, C.c_float(init_q), int(target),
                                              1 if cached else 0, C.c_float(cached.quality if cached else 0.0),
                                              C.c_double(cached.scale if cached else 1.0), out.ctypes.data, out.size,
                                              C.byref(t))
        if rc:
            raise self._err(rc)
        self._inflight[t.value] = (None, out)
        return t.value

    def submit_resize_image(self, img_handle, dw: int, dh: int, dst: Optional[np.ndarray]) -> int:
        """dst=None leaves the resized pixels in HBM (no device-to-host copy)."""
        t = C.c_int64(0)
        rc = self._L.icb_submit_resize_image(self._h, img_handle, dw, dh, dst.ctypes.data if dst is not None else None,
                                             dst.strides[0] if dst is not None else 0, C.byref(t))
        if rc:
            raise self._err(rc)
        self._inflight[t.value] = (None, dst)
        return t.value

    def submit_resize(self, image: np.ndarray, dst: np.ndarray, fmt: Optional[int] = None) -> int:
        """tools/ImageTools.java:7-26 in the queue: host pixels in, dst (dh, dw, c) filled when wait() returns."""
        image, w, h, pitch, fmt = _image_args(image, fmt)
        t = C.c_int64(0)
        rc = self._L.icb_submit_resize(self._h, image.ctypes.data, w, h, pitch, fmt, dst.ctypes.data, dst.shape[1],
                                       dst.shape[0], dst.strides[0], C.byref(t))
        if rc:
            raise self._err(rc)
        self._inflight[t.value] = (image, dst)
        return t.value

    def submit_png_fit(self, image: np.ndarray, min_w: int, min_h: int, dst: np.ndarray, fmt: Optional[int] = None):
        """core/ImageCompressionPng.java:37-66 in the queue -> (ticket, dw, dh), or None when the image is
        already inside the box (the seam returns false; nothing is queued).  dst: flat uint8 buffer."""
        image, w, h, pitch, fmt = _image_args(image, fmt)
        t, dw, dh = C.c_int64(0), C.c_int(0), C.c_int(0)
        rc = self._L.icb_submit_png_fit(self._h, image.ctypes.data, w, h, pitch, fmt, min_w, min_h, dst.ctypes.data,
                                        dst.size, C.byref(dw), C.byref(dh), C.byref(t))
        if rc == ICB_NO_FIT:
            return None
        if rc:
            raise self._err(rc)
        self._inflight[t.value] = (image, dst)
        return t.value, dw.value, dh.value

    def poll(self, ticket: int) -> bool:
        """True once the job has finished (wait() then returns at once); never blocks."""
        d = C.c_int(0)
        rc = self._L.icb_poll(self._h, ticket, C.byref(d))
        if rc:
            raise self._err(rc)
        return bool(d.value)

    # ---- the JPEG reader of core/ImageCompression.java:119-155 on the GPU ----
    def jpg_decode(self, data, fmt: int = FMT_RGB8) -> np.ndarray:
        """ImageIO's reader.read for a baseline 4:2:0 file -> (h, w, 3) pixels; IcbError(ICB_ERR_UNSUPPORTED) for the
        files the caller keeps decoding on the host."""
        buf = data if isinstance(data, np.ndarray) else np.frombuffer(data, np.uint8)
        w, h = jpg_file_info(buf)
        out = np.empty((h, w, 3), np.uint8)
        ww, hh = C.c_int(0), C.c_int(0)
        rc = self._L.icb_jpg_decode(self._h, buf.ctypes.data, buf.size, fmt, out.ctypes.data, out.size, w * 3,
                                    C.byref(ww), C.byref(hh))
        if rc:
            raise self._err(rc)
        return out

    def submit_jpg_file_fit(self, data, init_q: float, target: int, out: np.ndarray,
                            cached: Optional[LearnedParams] = None) -> int:
        buf = data if isinstance(data, np.ndarray) else np.frombuffer(data, np.uint8)
        t, w, h = C.c_int64(0), C.c_int(0), C.c_int(0)
        rc = self._L.icb_submit_jpg_file_fit(self._h, buf.ctypes.data, buf.size, C.c_float(init_q), int(target),
                                             1 if cached else 0, C.c_float(cached.quality if cached else 0.0),
                                             C.c_double(cached.scale if cached else 1.0), out.ctypes.data, out.size,
                                             C.byref(w), C.byref(h), C.byref(t))
        if rc:
            raise self._err(rc)
        self._inflight[t.value] = (buf, out)
        return t.value

    def jpg_file_fit(self, data, init_q: float, target: int, cached: Optional[LearnedParams] = None) -> FitResult:
        """decodeImageWithSubsampling + compressJpgWithTargetSize on the file bytes; .data = the JFIF output"""
        out = np.empty(max(int(target), 1), np.uint8)
        r = self.wait(self.submit_jpg_file_fit(data, init_q, target, out, cached))
        r.data = bytes(out[: r.out_len]) if r.status == ICB_OK else None
        return r

    # ---- the batch executor in native code (CompressionBatch.java:64-108) ----
    def make_items(self, specs) -> ItemList:
        """specs: dicts with kind ("jpg_fit" | "jpg_fit_image" | "png_fit" | "png_compress" | "resize_image"),
        image (host ndarray, or a resident handle for the *_image kinds), out (uint8 ndarray or None) and, by kind,
        init_q / target / cached (LearnedParams) or min_w / min_h or dw / dh."""
        arr = (_Item * len(specs))()
        keep = []
        for it, sp in zip(arr, specs):
            kind = sp["kind"]
            it.kind = ITEM_KINDS[kind]
            if kind in ("jpg_fit_image", "resize_image"):
                h = sp["image"]
                it.img = getattr(h, "value", h)
            elif kind == "jpg_file_fit":
                buf = sp["file"] if isinstance(sp["file"], np.ndarray) else np.frombuffer(sp["file"], np.uint8)
                it.px, it.pitch = buf.ctypes.data, buf.size
                keep.append(buf)
            else:
                image, w, hh, pitch, fmt = _image_args(sp["image"], sp.get("fmt"))
                it.px, it.w, it.h, it.pitch, it.fmt = image.ctypes.data, w, hh, pitch, fmt
                keep.append(image)
            it.init_q = sp.get("init_q", 0.0)
            it.target = int(sp.get("target", 0))
            c = sp.get("cached")
            it.has_cached = 1 if c else 0
            it.cached_q = c.quality if c else 0.0
            it.cached_scale = c.scale if c else 1.0
            it.min_w, it.min_h = sp.get("min_w", 0), sp.get("min_h", 0)
            it.dw, it.dh = sp.get("dw", 0), sp.get("dh", 0)
            out = sp.get("out")
            if out is not None:
                it.out, it.out_cap = out.ctypes.data, out.size
                keep.append(out)
        return ItemList(arr, keep)

    def run_list(self, items: ItemList, threads: int = 2, inflight: int = 3) -> None:
        """Runs every item through the library's own worker threads (one atomic cursor, `inflight` jobs per thread
        in the GPU work queues); results are in items.results() afterwards."""
        rc = self._L.icb_run_list(self._h, items.arr, len(items.arr), threads, inflight)
        if rc:
            raise self._err(rc)

    def wait(self, ticket: int) -> FitResult:
        r = _Result()
        rc = self._L.icb_wait(self._h, ticket, C.byref(r))
        self._inflight.pop(ticket, None)
        if rc < 0:
            raise self._err(rc)
        return FitResult(rc, None, float(r.best_q), float(r.scale), bool(r.used_cache), int(r.n_passes),
                         int(r.device), int(r.out_len))

    # ---- resident images ----
    def upload_image(self, image: np.ndarray, device_index: int = 0, fmt: Optional[int] = None):
        image, w, h, pitch, fmt = _image_args(image, fmt)
        p = C.c_void_p()
        rc = self._L.icb_image_upload(self._h, device_index, image.ctypes.data, w, h, pitch, fmt, C.byref(p))
        if rc:
            raise self._err(rc)
        return p

    def free_image(self, img_handle):
        self._L.icb_image_free(self._h, img_handle)

    def profile_jpg_fit_image(self, img_handle, init_q: float, target: int, reps: int = 1):
        """-> dict(transform_ms, search_ms, total_ms, entropy_kernel_ms, entropy_kernel_launches)"""
        a, b, c, d, n = C.c_float(0), C.c_float(0), C.c_float(0), C.c_float(0), C.c_int(0)
        rc = self._L.icb_profile_jpg_fit_image(self._h, img_handle, C.c_float(init_q), int(target), reps,
                                               C.byref(a), C.byref(b), C.byref(c), C.byref(d), C.byref(n))
        if rc:
            raise self._err(rc)
        return dict(transform_ms=float(a.value), search_ms=float(b.value), total_ms=float(c.value),
                    entropy_kernel_ms=float(d.value), entropy_kernel_launches=int(n.value))

    def profile_transform_series(self, img_handles, reps: int = 4, list_q: float = 0.0):
        """-> (ms per transform launch, launches timed): the images' transforms back to back on one stream;
        list_q > 0: with the survivor lists for that quality, as inside a fit"""
        arr = (C.c_void_p * len(img_handles))(*[getattr(x, "value", x) for x in img_handles])
        ms, n = C.c_float(0), C.c_int(0)
        rc = self._L.icb_profile_transform_series(self._h, arr, len(img_handles), reps, C.c_float(list_q),
                                                  C.byref(ms), C.byref(n))
        if rc:
            raise self._err(rc)
        return float(ms.value), int(n.value)

    def mark(self, device_index: int, which: int):
        rc = self._L.icb_mark(self._h, device_index, which)
        if rc:
            raise self._err(rc)

    def elapsed_ms(self, device_index: int) -> float:
        ms = C.c_float(0)
        rc = self._L.icb_elapsed_ms(self._h, device_index, C.byref(ms))
        if rc:
            raise self._err(rc)
        return float(ms.value)

    def stats(self) -> Dict[str, float]:
        s = _Stats()
        rc = self._L.icb_stats(self._h, C.byref(s))
        if rc:
            raise self._err(rc)
        d = {k: getattr(s, k) for k, _ in _Stats._fields_}
        d["kernel_clock_ns"] = list(s.kernel_clock_ns)              # [transform, entropy passes, emit]
        d["kernel_clock_launches"] = list(s.kernel_clock_launches)
        return d
