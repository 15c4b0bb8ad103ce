"""Build recipe for libparticles_cuda.so (sm_100a only) and the headless driver.

nvcc cross-compiles without a GPU.  -fmad=false is part of the numerical contract: the
reference's Zig code never fuses a multiply with an add, so neither may we.
"""
from __future__ import annotations

import os
import shutil
import subprocess
import sys

PKG_DIR = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(PKG_DIR)
CSRC = os.path.join(PKG_DIR, "csrc")
INCLUDE = os.path.join(ROOT, "include")
# PS_LIB=<path> (developer hook): load / build another file instead, e.g. an instrumented build made with
# PS_NVCC_EXTRA="-DPS_TAIL_PROFILE"; the product is always libparticles_cuda.so beside this file.
LIB_PATH = os.environ.get("PS_LIB") or os.path.join(PKG_DIR, "libparticles_cuda.so")
HEADLESS_PATH = os.path.join(PKG_DIR, "ps_headless")

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-O3", "-lineinfo", "-std=c++17",
    "-fmad=false", "-prec-div=true", "-prec-sqrt=true", "-ftz=false",
    "-Xcompiler", "-fPIC,-fno-fast-math,-ffp-contract=off",
    "-I", INCLUDE, "-I", CSRC,
]


def _nvcc() -> str:
    for cand in (shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found: libparticles_cuda cannot be built (there is no CPU fallback)")


def _stale(target: str, sources: list[str]) -> bool:
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(s) > t for s in sources)


def _sources() -> list[str]:
    out = [os.path.join(INCLUDE, "particles_cuda.h"), os.path.abspath(__file__)]
    for f in sorted(os.listdir(CSRC)):
        if f.endswith((".cu", ".cuh", ".h")):
            out.append(os.path.join(CSRC, f))
    return out


def build_lib(force: bool = False, verbose: bool = False) -> str:
    srcs = _sources()
    if not force and not _stale(LIB_PATH, srcs):
        return LIB_PATH
    cus = [s for s in srcs if s.endswith(".cu")]
    cmd = [_nvcc(), *NVCC_FLAGS, *os.environ.get("PS_NVCC_EXTRA", "").split(), "-shared", "-o", LIB_PATH, *cus, "-lcudart"]
    if verbose:
        cmd.insert(1, "-Xptxas=-v")
        print(" ".join(cmd), file=sys.stderr)
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError("nvcc failed:\n" + res.stdout + res.stderr)
    if verbose:
        print(res.stdout + res.stderr, file=sys.stderr)
    return LIB_PATH


def build_headless(force: bool = False) -> str:
    """The C++ replay of main.zig's call sequence (host/headless_main.cpp) linked against the library."""
    src = os.path.join(PKG_DIR, "host", "headless_main.cpp")
    hdr = os.path.join(PKG_DIR, "host", "core_cuda.hpp")
    if not os.path.exists(src):
        return ""
    build_lib(force)
    if not force and not _stale(HEADLESS_PATH, [src, hdr, LIB_PATH]):
        return HEADLESS_PATH
    cxx = os.environ.get("CXX") or shutil.which("g++") or "g++"
    cmd = [cxx, "-O2", "-std=c++17", "-ffp-contract=off", "-I", INCLUDE, "-I", os.path.join(PKG_DIR, "host"),
           src, "-o", HEADLESS_PATH, "-L", PKG_DIR, "-lparticles_cuda", "-Wl,-rpath,$ORIGIN"]
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError("headless build failed:\n" + res.stdout + res.stderr)
    return HEADLESS_PATH


if __name__ == "__main__":
    build_lib(force="--force" in sys.argv, verbose="-v" in sys.argv)
    build_headless(force="--force" in sys.argv)
    print(LIB_PATH)
