"""icb_png_encode_device_noalloc (no counterpart in the reference; include/imgcomp_b200.h): the PNG encoder on
device buffers with its work arena in a static __device__ array -- no cudaMalloc on any call.  Its file must be
byte-identical to the one icb_png_encode (the existing entry point, same kernels on a cudaMalloc'ed arena) makes
from the same pixels, on two calls in a row that share the arena.  Collected last on purpose
(file name)."""
import ctypes as C

import numpy as np
import pytest

import image_compression_b200 as pkg
from image_compression_b200 import api, synth

pytestmark = pytest.mark.gpu


def _noalloc(L, h, torch, img, fmt):
    hgt, w = img.shape[:2]
    bpp = 1 if img.ndim == 2 else img.shape[2]
    pitch = w * bpp + 5                                       # a ragged device pitch on purpose
    host = np.zeros((hgt, pitch), np.uint8)
    host[:, : w * bpp] = img.reshape(hgt, w * bpp)
    d_px = torch.from_numpy(host).cuda()
    cap = img.size * 2 + 65536
    d_out = torch.full((cap,), 0xAA, dtype=torch.uint8, device="cuda")
    d_len = torch.zeros(1, dtype=torch.int64, device="cuda")
    torch.cuda.synchronize()
    rc = L.icb_png_encode_device_noalloc(h, d_px.data_ptr(), w, hgt, pitch, fmt, d_out.data_ptr(), cap,
                                         d_len.data_ptr(), None)
    assert rc == api.ICB_OK, L.icb_last_error(h)
    torch.cuda.synchronize()
    n = int(d_len.cpu()[0])
    assert 0 < n <= cap
    return d_out[:n].cpu().numpy().tobytes()


def test_noalloc_png_equals_the_allocating_entry_point_twice():
    import torch
    L = pkg.load_library()
    with pkg.Engine(devices=[0], slots_per_gpu=1) as eng:
        torch.cuda.set_device(0)
        cases = [(synth.natural_image(21, 300, 217), api.FMT_RGB8),          # 2nd call: smaller image, other format,
                 (np.dstack([synth.natural_image(22, 190, 133),              # stale arena contents behind it
                             synth.natural_image(23, 190, 133)[:, :, :1]]), api.FMT_RGBA8),
                 (synth.natural_image(24, 64, 40)[:, :, 0].copy(), api.FMT_GRAY8)]
        for img, fmt in cases:
            want = eng.png_encode(np.ascontiguousarray(img), fmt)
            got = _noalloc(L, eng._h, torch, np.ascontiguousarray(img), fmt)
            assert got == want, (img.shape, fmt, len(got), len(want))
        # repeat the first case: the arena now holds the third call's leftovers
        img, fmt = cases[0]
        assert _noalloc(L, eng._h, torch, img, fmt) == eng.png_encode(img, fmt)
        # handle-less call, and the size limit
        assert _noalloc(L, None, torch, img, fmt) == eng.png_encode(img, fmt)
        big = torch.zeros(1024 * 1024 * 3 + 16, dtype=torch.uint8, device="cuda")
        out = torch.zeros(1 << 20, dtype=torch.uint8, device="cuda")
        n = torch.zeros(1, dtype=torch.int64, device="cuda")
        rc = L.icb_png_encode_device_noalloc(eng._h, big.data_ptr(), 1024, 1024, 3072, api.FMT_RGB8, out.data_ptr(),
                                             1 << 20, n.data_ptr(), None)
        assert rc == api.ICB_ERR_CAPACITY
        assert b"exceeds the static arena" in L.icb_last_error(eng._h)
        torch.cuda.synchronize()
