"""CPU suite: the decode oracle (oracle/jpeg_decode_oracle.c) against the committed
Pillow / libjpeg-turbo fixtures (tests/golden/make_golden_decode.py) -- bit-exact pixels for baseline 4:2:0,
4:2:2 and 4:4:4 files, and a refusal for everything the CUDA decoder does not take."""
import os

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
Z = np.load(os.path.join(HERE, "golden", "jpeg_decode.npz"))
N = len([k for k in Z.files if k.startswith("file_")])


@pytest.mark.parametrize("i", range(N))
def test_oracle_decode_equals_libjpeg_turbo(oracle, i):
    data, want = Z[f"file_{i}"].tobytes(), Z[f"pixels_{i}"]
    assert oracle.jpeg_info(data) == (want.shape[1], want.shape[0])
    assert np.array_equal(oracle.jpeg_decode(data), want)
    assert np.array_equal(oracle.jpeg_decode(data, oracle.FMT_BGR8), want[:, :, ::-1])


def test_coefficients_round_trip_through_the_forward_oracle(oracle):
    """decode(encode(img)) coefficients = quantised forward coefficients: ties the decode oracle to the encode
    oracle (same block order, DC chains, zigzag), independent of the fixtures."""
    from image_compression_b200 import synth
    img = synth.natural_image(77, 150, 90)
    lum, chr_ = oracle.qtables(0.6)
    data = oracle.encode(img, 0.6)
    got = oracle.jpeg_decode_coefficients(data)                    # natural order, quantised
    fwd = oracle.forward_coefficients(img).reshape(-1, 64)         # zigzag order, unquantised (8 x DCT)
    zz = np.array([0, 1, 8, 16, 9, 2, 3, 10, 17, 24, 32, 25, 18, 11, 4, 5, 12, 19, 26, 33, 40, 48, 41, 34, 27, 20, 13, 6, 7,
                   14, 21, 28, 35, 42, 49, 56, 57, 50, 43, 36, 29, 22, 15, 23, 30, 37, 44, 51, 58, 59, 52, 45, 38, 31, 39, 46,
                   53, 60, 61, 54, 47, 55, 62, 63])
    nat = np.empty_like(fwd)
    nat[:, zz] = fwd
    for b in range(nat.shape[0]):
        t = (lum if b % 6 < 4 else chr_).astype(np.int64).reshape(64) * 8
        x = nat[b].astype(np.int64)
        q = np.sign(x) * ((np.abs(x) + t // 2) // t)
        assert np.array_equal(q, got[b].astype(np.int64)), b


@pytest.mark.parametrize("name", ["progressive", "gray"])
def test_oracle_refuses_what_the_gpu_decoder_does_not_take(oracle, name):
    with pytest.raises(ValueError) as ei:
        oracle.jpeg_info(Z["unsup_" + name].tobytes())
    assert ei.value.args[0] == oracle.DEC_UNSUPPORTED


def test_truncated_and_garbage_files_do_not_crash(oracle):
    data = Z["file_2"].tobytes()
    for cut in (0, 1, 3, 20, 200, len(data) // 2, len(data) - 2):
        try:
            oracle.jpeg_decode(data[:cut])
        except ValueError:
            pass
    rng = np.random.default_rng(5)
    for _ in range(20):
        junk = bytearray(data)
        for k in rng.integers(700, len(junk), 8):
            junk[k] = int(rng.integers(0, 256))
        try:
            oracle.jpeg_decode(bytes(junk))
        except ValueError:
            pass


@pytest.mark.parametrize("variant", ["extra_segments", "merged_tables", "fill_and_tail"])
def test_rewrapped_files_decode_alike(oracle, variant):
    """Exif / comment segments, merged table segments, fill bytes, bytes behind EOI (tests/jpeg_mangle.py)"""
    import sys
    sys.path.insert(0, HERE)
    import jpeg_mangle
    for i in (2, 13, 19, 24):                                   # 4:2:0, 4:2:2, 4:4:4, restart intervals
        data = jpeg_mangle.VARIANTS[variant](Z[f"file_{i}"].tobytes())
        assert np.array_equal(oracle.jpeg_decode(data), Z[f"pixels_{i}"]), (variant, i)
