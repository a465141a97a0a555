"""CPU check of the transform kernel's arithmetic forms (csrc/xf_arith.cuh: dp2a colour conversion on
packed pixels, row pass on sample bytes, column pass with doubled constants) against the CPU oracle.
The header compiles for the host (dp2a / dp4a / prmt emulated as the PTX ISA defines them), so the
exact-integer identities the kernel relies on are pinned here without a GPU; the GPU parity tests then
cover the kernel's tiling, edges and dummy blocks."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def host_bin(tmp_path_factory):
    from oracle import oracle as O
    O.build()
    out = str(tmp_path_factory.mktemp("xf") / "xf_arith_host")
    odir = os.path.join(ROOT, "oracle")
    cmd = ["g++", "-O2", "-std=c++17", "-o", out, os.path.join(ROOT, "tests", "host", "xf_arith_host.cpp"),
           "-L" + odir, "-loracle", "-Wl,-rpath," + odir]
    subprocess.run(cmd, check=True, capture_output=True, text=True)
    return out


@pytest.mark.parametrize("w,h,seed,mode", [(64, 48, 1, 0), (128, 64, 2, 1), (256, 32, 3, 2), (16, 16, 4, 0),
                                           (512, 256, 5, 0), (512, 256, 6, 1)])
def test_transform_arithmetic_matches_oracle(host_bin, w, h, seed, mode):
    r = subprocess.run([host_bin, str(w), str(h), str(seed), str(mode)], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "mismatches 0" in r.stdout
