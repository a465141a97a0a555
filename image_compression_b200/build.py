"""Builds libimgcomp_b200.so in-tree with nvcc for sm_100a (no JIT cache: the .so
travels with the repo snapshot to the GPU box)."""
from __future__ import annotations

import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libimgcomp_b200.so")
SOURCES = ["api.cu"]
DEPS = ["api.cu", "common.cuh", "xf_arith.cuh", "k_transform.cuh", "k_entropy.cuh", "k_control.cuh", "k_resize.cuh",
        "k_png.cuh", "k_decode.cuh", "jpeg_parse.h",
        os.path.join("..", "..", "include", "imgcomp_b200.h")]


def nvcc_path() -> str:
    for cand in (os.environ.get("NVCC"), "/usr/local/cuda/bin/nvcc", shutil.which("nvcc")):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found")


def needs_build() -> bool:
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    return any(os.path.getmtime(os.path.join(CSRC, d)) > t for d in DEPS)


def build(force: bool = False, verbose: bool = False) -> str:
    if not force and not needs_build():
        return LIB
    cmd = [nvcc_path(), "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17",
           "--expt-relaxed-constexpr", "-Xcompiler", "-fPIC,-O3,-fno-fast-math", "-shared", "-cudart", "static",
           "-o", LIB] + [os.path.join(CSRC, s) for s in SOURCES]
    if os.environ.get("ICB_TIMING"):
        cmd.insert(1, "-DICB_TIMING")
    for knob in ("ICB_TILE_MCUS", "ICB_XF_CTAS", "ICB_XF_STAGES", "ICB_PASS_LDG", "ICB_NO_KCLK", "ICB_PASS_BITS_ONLY", "ICB_XF_ROWMAJOR", "ICB_XF_DBG", "ICB_XF_COLEXPANDED", "ICB_PASS_OR_BRANCH", "ICB_DEC_THREADS"):        # A/B builds of the transform kernel
        if os.environ.get(knob):
            cmd.insert(1, "-D%s=%s" % (knob, os.environ[knob]))
    if os.environ.get("ICB_LIB_SUFFIX"):
        cmd[cmd.index("-o") + 1] = LIB.replace(".so", os.environ["ICB_LIB_SUFFIX"] + ".so")
    if verbose:
        cmd.insert(1, "-Xptxas")
        cmd.insert(2, "-v")
    r = subprocess.run(cmd, capture_output=True, text=True)
    if verbose or r.returncode:
        sys.stderr.write(r.stdout + r.stderr)
    if r.returncode:
        raise RuntimeError("nvcc failed")
    return LIB


if __name__ == "__main__":
    print(build(force=True, verbose="-v" in sys.argv))
