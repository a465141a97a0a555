"""CPU suite, part 2: the C-ABI library loads and exports every symbol include/*.h declares,
host-only entry points answer without a GPU, and compute entry points fail loudly (there is
no CPU fallback to route through).  No compute calls are made here."""
import ctypes as C
import os
import re

import numpy as np
import pytest

import image_compression_b200 as pkg
from image_compression_b200 import api

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    syms = set()
    inc = os.path.join(ROOT, "include")
    for fn in os.listdir(inc):
        if fn.endswith(".h"):
            text = open(os.path.join(inc, fn)).read()
            text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
            syms.update(re.findall(r"\b(icb_[a-z0-9_]+)\s*\(", text))
    return syms


def test_header_and_binding_agree():
    declared = _declared_symbols()
    assert declared == set(api.ABI_SYMBOLS), declared ^ set(api.ABI_SYMBOLS)


def test_library_exports_every_declared_symbol():
    L = pkg.load_library()
    assert os.path.samefile(pkg.library_path(), os.path.join(ROOT, "image_compression_b200", "libimgcomp_b200.so"))
    for s in sorted(_declared_symbols()):
        assert getattr(L, s) is not None, s


def test_status_codes_match_header():
    text = open(os.path.join(ROOT, "include", "imgcomp_b200.h")).read()
    codes = dict(re.findall(r"(ICB_[A-Z0-9_]+)\s*=\s*(-?\d+)", text))
    assert int(codes["ICB_OK"]) == api.ICB_OK and int(codes["ICB_NO_FIT"]) == api.ICB_NO_FIT
    assert int(codes["ICB_ERR_INVALID"]) == api.ICB_ERR_INVALID and int(codes["ICB_ERR_OOM"]) == api.ICB_ERR_OOM
    assert int(codes["ICB_ERR_CAPACITY"]) == api.ICB_ERR_CAPACITY
    assert int(codes["ICB_ERR_NO_DEVICE"]) == api.ICB_ERR_NO_DEVICE
    assert int(codes["ICB_FMT_RGB8"]) == api.FMT_RGB8 and int(codes["ICB_FMT_ABGR8"]) == api.FMT_ABGR8


def test_result_and_stats_struct_layout():
    # icb_result / icb_stats_t as the header lays them out on LP64
    assert C.sizeof(api._Result) == 48
    assert api._Result.out_len.offset == 8 and api._Result.scale.offset == 24 and api._Result.device.offset == 40
    assert C.sizeof(api._Stats) == 112 and api._Stats.kernel_clock_ns.offset == 64
    assert api._Stats.last_transform_ms.offset == 40


def test_host_only_entry_points_need_no_device():
    # tools/ImageTools.java:8-9 and core/ImageCompressionPng.java:49-61
    assert pkg.resize_dims(800, 600, 0.125) == (100, 75)
    assert pkg.resize_dims(4000, 3000, 0.85) == (3400, 2550)
    assert pkg.resize_dims(5, 5, 0.01) == (1, 1)
    assert pkg.png_target_dims(100, 100, 100, 100) is None
    assert pkg.png_target_dims(800, 600, 100, 100)[:2] == (100, 75)
    assert pkg.png_target_dims(7680, 4320, 1920, 1920) == (1920, 1080, 0.25)
    with pytest.raises(pkg.IcbError):
        pkg.png_target_dims(0, 10, 5, 5)
    # core/ImageCompression.java:140-153: highestOneBit(floor(maxDim / 4096)) above 4096
    for (w, h), s in {(4000, 3000): 1, (4096, 4096): 1, (6000, 4000): 1, (8191, 100): 1, (8192, 6000): 2,
                      (11548, 8660): 2, (12288, 100): 2, (16384, 16384): 4, (30000, 20000): 4, (40000, 1): 8}.items():
        assert pkg.subsample_factor(w, h) == s, (w, h)


def test_null_handle_is_an_error_not_a_crash():
    L = pkg.load_library()
    n = C.c_size_t(0)
    buf = np.zeros(16, np.uint8)
    assert L.icb_jpg_encode(None, buf.ctypes.data, 2, 2, 6, 0, C.c_float(0.5), buf.ctypes.data, 16, C.byref(n)) == \
        api.ICB_ERR_INVALID
    assert L.icb_wait(None, 0, None) == api.ICB_ERR_INVALID
    assert L.icb_device_count(None) == 0
    L.icb_shutdown(None)
    L.icb_free_pinned(None)


def _has_cuda():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


@pytest.mark.skipif(_has_cuda(), reason="only meaningful on a box without a GPU")
def test_no_cpu_fallback_without_a_device():
    with pytest.raises(pkg.IcbError) as ei:
        pkg.Engine()
    assert ei.value.code == api.ICB_ERR_NO_DEVICE


def test_product_package_never_touches_the_oracle():
    """oracle/ is test infrastructure: nothing under the package may import, load or link it."""
    pkg_dir = os.path.join(ROOT, "image_compression_b200")
    for dirpath, _, files in os.walk(pkg_dir):
        for fn in files:
            if fn.endswith((".py", ".cu", ".cuh", ".h", ".cpp")):
                text = open(os.path.join(dirpath, fn), errors="replace").read()
                assert "liboracle" not in text and "from oracle" not in text and "import oracle" not in text, fn


def test_jpeg_header_parse_needs_no_device():
    """icb_jpg_file_info (core/ImageCompression.java:128-129 reader.getWidth / getHeight): host arithmetic only"""
    z = np.load(os.path.join(ROOT, "tests", "golden", "jpeg_decode.npz"))
    n = len([k for k in z.files if k.startswith("file_")])
    for i in range(n):
        h, w = z[f"pixels_{i}"].shape[:2]
        assert pkg.jpg_file_info(z[f"file_{i}"]) == (w, h)
    for name in ("progressive", "gray"):
        with pytest.raises(pkg.IcbError) as ei:
            pkg.jpg_file_info(z["unsup_" + name])
        assert ei.value.code == api.ICB_ERR_UNSUPPORTED
    for junk in (b"", b"\xff\xd8", b"not a jpeg at all", z["file_0"].tobytes()[:100]):
        with pytest.raises(pkg.IcbError) as ei:
            pkg.jpg_file_info(junk if junk else np.zeros(0, np.uint8))
        assert ei.value.code == api.ICB_ERR_INVALID
