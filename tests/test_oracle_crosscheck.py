"""The compiled C++ oracle (oracle/core_ref.hpp) against the independent pure-Python restatement
(oracle/core_ref.py): bit-for-bit equal positions, velocities and ordered pair lists on small
random scenes.  Both were written separately from /root/reference/src/core.zig; agreement is the
second thing (after the hand-derived KATs) that pins the oracle — parity with the Zig binary
itself stays unpinned (no toolchain, no reference vectors)."""
import numpy as np
import pytest

import oracle
from oracle import core_ref as pyref
from particle_simulator_b200 import scenes


def python_run(scene, pos, vel, substeps, *, collisions, t_acc, real, capture=-1):
    par = pyref.Particles(pos, vel, scene.radius, real=real)
    sol = pyref.Solver(scene.target_fps, scene.sub_steps, scene.g, scene.boundary_damping, scene.particle_damping, real=real)
    sol.t_acc = t_acc
    shg = pyref.SpatialHashGrid(scene.width, scene.height, scene.grid_size, len(pos), real=real)
    pairs = None
    for s in range(substeps):
        sol.substep(par, scene.width, scene.height, shg, collisions=collisions)
        if s == capture:
            pairs = list(shg.collision_pairs)
    dt = np.float64 if real is float else np.float32
    return np.array(par.positions, dtype=dt), np.array(par.velocities, dtype=dt), pairs


def small_scene(n, seed, dense=True):
    # a small world at ~2x the reference density so that many contacts (and same-cell pairs) occur
    side = 4.0 * max(int((n / (0.9 if dense else 0.45)) ** 0.5), 3)
    return scenes.Scene(count=n, width=side, height=side + 8.0), seed


@pytest.mark.parametrize("real,dtype", [(float, np.float64), (np.float32, np.float32)])
@pytest.mark.parametrize("t_acc", [False, True])
@pytest.mark.parametrize("n,seed,substeps", [(150, 1, 24), (400, 2, 12)])
def test_cpp_oracle_equals_python_restatement(real, dtype, t_acc, n, seed, substeps):
    scene, seed = small_scene(n, seed)
    pos0, vel0 = scenes.spawn(scene, seed=seed, dtype=dtype)
    p = oracle.make_params(radius=scene.radius, width=scene.width, height=scene.height, grid_size=scene.grid_size)
    pos_c, vel_c = pos0.copy(), vel0.copy()
    out = oracle.run(p, pos_c, vel_c, substeps, collisions=True, t_acc=t_acc, capture_substep=substeps - 1, pairs_cap=8 * n)
    pos_p, vel_p, pairs_p = python_run(scene, pos0, vel0, substeps, collisions=True, t_acc=t_acc, real=real, capture=substeps - 1)
    assert out["stats"]["pairs_total"] > n // 4, "scene too sparse to exercise the collision path"
    assert out["pairs"].tolist() == [list(x) for x in pairs_p]
    assert np.array_equal(pos_c.view(np.uint8), pos_p.view(np.uint8))
    assert np.array_equal(vel_c.view(np.uint8), vel_p.view(np.uint8))


@pytest.mark.parametrize("real,dtype", [(float, np.float64), (np.float32, np.float32)])
def test_integrate_only_equals_python_restatement(real, dtype):
    scene = scenes.Scene(count=500, width=200.0, height=120.0)
    pos0, vel0 = scenes.spawn(scene, seed=3, dtype=dtype)
    p = oracle.make_params(width=200.0, height=120.0)
    pos_c, vel_c = pos0.copy(), vel0.copy()
    oracle.run(p, pos_c, vel_c, 40, collisions=False, t_acc=True)
    pos_p, vel_p, _ = python_run(scene, pos0, vel0, 40, collisions=False, t_acc=True, real=real)
    assert np.array_equal(pos_c.view(np.uint8), pos_p.view(np.uint8))
    assert np.array_equal(vel_c.view(np.uint8), vel_p.view(np.uint8))


def test_vector_boundary_matches_the_unused_scalar_variant():
    # src/core.zig:269-288 (dead code) and :290-318 agree whenever a component violates one wall only
    scene = scenes.Scene(count=300, width=100.0, height=80.0)
    pos, vel = scenes.spawn(scene, seed=5)
    pos += vel * 0.02  # push many particles through the walls
    a = pyref.Particles(pos, vel, 2.0)
    b = pyref.Particles(pos, vel, 2.0)
    sol = pyref.Solver(60.0, 8, 50.0, 0.1, 0.9)
    sol._apply_boundary(a, 100.0, 80.0)
    sol.apply_boundary_scalar(b, 100.0, 80.0)
    assert a.positions == b.positions and a.velocities == b.velocities
