"""Shared helpers for the GPU parity tests: run the same seeded scene through libparticles_cuda
(via the C ABI) and through the oracle, and compare bit patterns."""
from __future__ import annotations

import numpy as np

import oracle
from particle_simulator_b200 import Simulation, scenes

DTYPES = {"f64": np.float64, "f32": np.float32}


def oracle_params(scene: scenes.Scene):
    return oracle.make_params(radius=scene.radius, width=scene.width, height=scene.height, grid_size=scene.grid_size,
                              target_fps=scene.target_fps, sub_steps=scene.sub_steps, g=scene.g,
                              boundary_damping=scene.boundary_damping, particle_damping=scene.particle_damping)


def run_oracle(scene, pos0, vel0, substeps, *, collisions, t_acc=False, capture=-1, pairs_cap=0):
    pos, vel = pos0.copy(), vel0.copy()
    out = oracle.run(oracle_params(scene), pos, vel, substeps, collisions=collisions, t_acc=t_acc,
                     capture_substep=capture, pairs_cap=pairs_cap)
    return pos, vel, out


def run_gpu(scene, pos0, vel0, substeps, *, precision, collisions, t_acc=False, capture_last=False, **kw):
    """Upload -> step -> download through the C ABI.  Returns pos, vel (dtype of pos0), pairs, stats."""
    with Simulation(scene, precision=precision, collisions=collisions, **kw) as sim:
        sim.upload(pos0, vel0)
        sim.set_acc(t_acc)
        pairs = None
        if capture_last and collisions:
            sim.capture_pairs(True)
        sim.step_substeps(substeps)
        pos, vel = sim.download(dtype=pos0.dtype)
        if capture_last and collisions:
            pairs = sim.contact_pairs()
        stats = sim.stats()
    return pos, vel, pairs, stats


def bits_equal(a: np.ndarray, b: np.ndarray) -> bool:
    return a.dtype == b.dtype and a.shape == b.shape and np.array_equal(a.view(np.uint8), b.view(np.uint8))


def max_ulp(a: np.ndarray, b: np.ndarray) -> int:
    it = np.int64 if a.dtype == np.float64 else np.int32
    ia, ib = a.view(it).astype(np.int64), b.view(it).astype(np.int64)
    return int(np.abs(ia - ib).max()) if a.size else 0


def assert_bit_exact(got, want, what=""):
    assert got.shape == want.shape and got.dtype == want.dtype
    if not bits_equal(got, want):
        bad = np.nonzero((got.view(np.uint8) != want.view(np.uint8)).reshape(got.shape[0], -1).any(axis=1))[0]
        rel = np.abs(got - want).max() / max(np.abs(want).max(), 1e-300)
        raise AssertionError(f"{what}: {bad.size} of {got.shape[0]} particles differ (first {bad[:5].tolist()}), "
                             f"max ulp {max_ulp(got, want)}, max rel {rel:.3e}")
