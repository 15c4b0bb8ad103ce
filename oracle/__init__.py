"""oracle — CPU restatement of /root/reference/src/core.zig's particle step.

TEST INFRASTRUCTURE ONLY.  PARITY UNPINNED (see core_ref.hpp).  Importable only from
tests/, ``__graft_entry__.smoke()`` and bench.py's cpu_baseline / ``--impl reference`` legs.

``oracle.run(...)`` drives the compiled C++ restatement (liboracle_core.so, built by
``make -C oracle``) through ctypes; ``oracle.core_ref`` is the independent pure-Python one.
"""
from __future__ import annotations

import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "liboracle_core.so")
_lib = None


class Params(ctypes.Structure):
    """Mirror of oracle::Params (core_ref.hpp)."""

    _fields_ = [
        ("radius", ctypes.c_double),
        ("width", ctypes.c_double),
        ("height", ctypes.c_double),
        ("grid_size", ctypes.c_double),
        ("target_fps", ctypes.c_double),
        ("sub_steps", ctypes.c_uint64),
        ("g", ctypes.c_double),
        ("boundary_damping", ctypes.c_double),
        ("particle_damping", ctypes.c_double),
    ]


class Stats(ctypes.Structure):
    _fields_ = [
        ("pairs_total", ctypes.c_uint64),
        ("pairs_last", ctypes.c_uint64),
        ("max_cell_count", ctypes.c_uint32),
        ("cell_overflow", ctypes.c_uint32),
    ]


def build(force: bool = False) -> str:
    """Compile liboracle_core.so with oracle/Makefile (g++ -O2 -ffp-contract=off)."""
    srcs = [os.path.join(_HERE, f) for f in ("oracle_capi.cpp", "core_ref.hpp", "Makefile")]
    stale = force or not os.path.exists(_LIB_PATH) or any(
        os.path.getmtime(s) > os.path.getmtime(_LIB_PATH) for s in srcs
    )
    if stale:
        subprocess.run(["make", "-C", _HERE, "-B" if force else "-s"], check=True, capture_output=True)
    return _LIB_PATH


def lib():
    global _lib
    if _lib is None:
        build()
        L = ctypes.CDLL(_LIB_PATH)
        for name, real in (("oracle_run_f64", ctypes.c_double), ("oracle_run_f32", ctypes.c_float)):
            fn = getattr(L, name)
            rp = ctypes.POINTER(real)
            fn.argtypes = [
                ctypes.POINTER(Params), ctypes.c_uint64, rp, rp, rp,
                ctypes.c_uint64, ctypes.c_int, ctypes.c_int,
                ctypes.c_int64, ctypes.POINTER(ctypes.c_uint64), ctypes.c_uint64, ctypes.POINTER(ctypes.c_uint64),
                ctypes.POINTER(Stats), ctypes.POINTER(ctypes.c_double),
            ]
            fn.restype = ctypes.c_int
        L.oracle_grid_shape.argtypes = [ctypes.POINTER(Params), ctypes.POINTER(ctypes.c_uint64), ctypes.POINTER(ctypes.c_uint64)]
        L.oracle_grid_shape.restype = ctypes.c_int
        L.oracle_dt_f64.argtypes = [ctypes.POINTER(Params)]
        L.oracle_dt_f64.restype = ctypes.c_double
        L.oracle_dt_f32.argtypes = [ctypes.POINTER(Params)]
        L.oracle_dt_f32.restype = ctypes.c_float
        _lib = L
    return _lib


def make_params(radius=2.0, width=720.0, height=640.0, grid_size=4.0, target_fps=60.0, sub_steps=8,
                g=50.0, boundary_damping=0.1, particle_damping=0.9) -> Params:
    """Defaults = the main.zig scene (src/main.zig:25-29,63-75)."""
    return Params(radius, width, height, grid_size, target_fps, sub_steps, g, boundary_damping, particle_damping)


def run(params: Params, pos: np.ndarray, vel: np.ndarray, substeps: int, *, collisions: bool, t_acc: bool = False,
        acc: np.ndarray | None = None, capture_substep: int = -1, pairs_cap: int = 0):
    """Advance `substeps` sub-steps in place on interleaved (n,2) arrays (float64 or float32).

    Returns dict(pairs=(m,2) uint64 array of the captured sub-step or None, pairs_count, stats, seconds).
    """
    assert pos.dtype == vel.dtype and pos.dtype in (np.float64, np.float32)
    assert pos.flags.c_contiguous and vel.flags.c_contiguous and pos.shape == vel.shape
    n = pos.shape[0]
    if acc is None:
        acc = np.zeros_like(pos)
    L = lib()
    is64 = pos.dtype == np.float64
    fn = L.oracle_run_f64 if is64 else L.oracle_run_f32
    real = ctypes.c_double if is64 else ctypes.c_float
    rp = ctypes.POINTER(real)
    pairs = np.zeros((max(pairs_cap, 1), 2), dtype=np.uint64)
    count = ctypes.c_uint64(0)
    stats = Stats()
    secs = ctypes.c_double(0)
    rc = fn(ctypes.byref(params), n, pos.ctypes.data_as(rp), vel.ctypes.data_as(rp), acc.ctypes.data_as(rp),
            substeps, int(collisions), int(t_acc), capture_substep,
            pairs.ctypes.data_as(ctypes.POINTER(ctypes.c_uint64)), pairs_cap, ctypes.byref(count),
            ctypes.byref(stats), ctypes.byref(secs))
    if rc != 0:
        raise RuntimeError(f"oracle_run failed: {rc}")
    m = min(int(count.value), pairs_cap)
    return {
        "pairs": pairs[:m].copy() if capture_substep >= 0 else None,
        "pairs_count": int(count.value),
        "stats": {"pairs_total": int(stats.pairs_total), "pairs_last": int(stats.pairs_last),
                  "max_cell_count": int(stats.max_cell_count), "cell_overflow": int(stats.cell_overflow)},
        "seconds": float(secs.value),
        "acc": acc,
    }


def grid_shape(params: Params):
    r, c = ctypes.c_uint64(0), ctypes.c_uint64(0)
    if lib().oracle_grid_shape(ctypes.byref(params), ctypes.byref(r), ctypes.byref(c)) != 0:
        raise RuntimeError("bad grid parameters")
    return int(r.value), int(c.value)
