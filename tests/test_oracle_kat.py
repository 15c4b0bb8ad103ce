"""Known-answer tests for the oracle, written by hand from the formulas in
/root/reference/src/core.zig (the reference ships no tests or golden vectors, so these and
the independent Python restatement are what pins the oracle — "parity unpinned").

Every expected value below is computed in the test from the reference's formula with plain
Python float (IEEE f64) arithmetic, one rounded operation at a time; none of it calls the oracle.
"""
import numpy as np
import pytest

import oracle

DT = (1.0 / 60.0) / 8.0  # Solver.init, src/core.zig:227-228 with main.zig's fps=60, sub_steps=8


def run1(pos, vel, *, collisions, t_acc=False, substeps=1, capture=-1, dtype=np.float64, **kw):
    p = oracle.make_params(**kw)
    pos = np.array(pos, dtype=dtype).reshape(-1, 2).copy()
    vel = np.array(vel, dtype=dtype).reshape(-1, 2).copy()
    out = oracle.run(p, pos, vel, substeps, collisions=collisions, t_acc=t_acc, capture_substep=capture, pairs_cap=64)
    return pos, vel, out


def test_dt_derivation():
    p = oracle.make_params()
    assert oracle.lib().oracle_dt_f64(p) == DT
    assert oracle.lib().oracle_dt_f32(p) == np.float32(np.float32(1) / np.float32(60)) / np.float32(8)


def test_grid_shape_default_and_truncation():
    assert oracle.grid_shape(oracle.make_params()) == (160, 180)  # rows, cols  (src/core.zig:66-67)
    # non-integer sizes are truncated before the integer division (src/core.zig:61-63)
    assert oracle.grid_shape(oracle.make_params(width=721.9, height=643.9, grid_size=4.7)) == (160, 180)


def test_free_flight_no_gravity():
    # vel += 0*dt ; pos += vel*dt   (src/core.zig:252-259)
    pos, vel, _ = run1([100.0, 200.0], [48.0, -96.0], collisions=False)
    assert pos[0, 0] == 100.0 + 48.0 * DT
    assert pos[0, 1] == 200.0 + (-96.0) * DT
    assert vel[0, 0] == 48.0 and vel[0, 1] == -96.0


def test_gravity_semi_implicit_euler():
    # acc = (acc.x, g); NEW velocity is used for the position update (src/core.zig:249-259)
    pos, vel, out = run1([100.0, 200.0], [0.0, 0.0], collisions=False, t_acc=True, substeps=2)
    v1 = 0.0 + 50.0 * DT
    y1 = 200.0 + v1 * DT
    v2 = v1 + 50.0 * DT
    y2 = y1 + v2 * DT
    assert vel[0, 1] == v2 and pos[0, 1] == y2
    assert vel[0, 0] == 0.0 and pos[0, 0] == 100.0
    assert out["acc"][0, 0] == 0.0 and out["acc"][0, 1] == 50.0


def test_negative_zero_velocity_becomes_positive_zero():
    # -0.0 + 0.0*dt = +0.0: the "+ acc*dt" must not be dropped when acc = 0
    pos, vel, _ = run1([100.0, 100.0], [-0.0, -0.0], collisions=False)
    assert not np.signbit(vel[0, 0]) and not np.signbit(vel[0, 1])


@pytest.mark.parametrize("axis", [0, 1])
def test_wall_low(axis):
    # pos < r: vel *= -boundary_damping (restitution 0.1, Q7), pos clamped to r (src/core.zig:301-316)
    pos0 = [100.0, 100.0]; vel0 = [0.0, 0.0]
    pos0[axis] = 3.0; vel0[axis] = -600.0
    pos, vel, _ = run1(pos0, vel0, collisions=False)
    assert 3.0 + (-600.0) * DT < 2.0
    assert pos[0, axis] == 2.0
    assert vel[0, axis] == -600.0 * -0.1
    assert pos[0, 1 - axis] == 100.0 and vel[0, 1 - axis] == 0.0


@pytest.mark.parametrize("axis,dim", [(0, 720.0), (1, 640.0)])
def test_wall_high(axis, dim):
    pos0 = [100.0, 100.0]; vel0 = [0.0, 0.0]
    pos0[axis] = dim - 3.0; vel0[axis] = 600.0
    pos, vel, _ = run1(pos0, vel0, collisions=False)
    assert pos[0, axis] == dim - 2.0
    assert vel[0, axis] == 600.0 * -0.1


def test_wall_exact_boundary_is_not_reflected():
    # comparisons are strict (src/core.zig:306-307): a particle exactly at r or dim-r keeps its velocity
    pos, vel, _ = run1([2.0, 638.0], [0.0, 0.0], collisions=False)
    assert pos[0, 0] == 2.0 and pos[0, 1] == 638.0
    pos, vel, _ = run1([2.0 + 480.0 * DT, 300.0], [-480.0, 0.0], collisions=False)
    x = (2.0 + 480.0 * DT) + (-480.0) * DT
    if x == 2.0:
        assert vel[0, 0] == -480.0


def test_spawn_outside_wall_is_clamped_and_reflected_once():
    # Q8: spawn range is [r, W) so x may start beyond W - r; one scaling even though the mask is an OR
    pos, vel, _ = run1([719.5, 100.0], [10.0, 0.0], collisions=False)
    assert pos[0, 0] == 718.0 and vel[0, 0] == 10.0 * -0.1


def test_same_cell_pair_listed_twice_and_position_correction():
    # Two resting particles in one cell, 2 px apart (overlap 2).  Q1: listed twice; Q10: the second
    # entry was detected with the frozen positions but resolves with overlap = 0 (no correction).
    pos, vel, out = run1([101.0, 101.0, 103.0, 101.0], [0.0, 0.0, 0.0, 0.0], collisions=True, capture=0)
    assert out["pairs"].tolist() == [[1, 0], [0, 1]]          # [neighbour, centre] (Q4), centres ascending
    # entry [i=1, j=0]: n = (p0 - p1)/2 = (-1, 0); correction = overlap*0.5 = 1; p1 -= 1*n ; p0 += 1*n
    assert pos.tolist() == [[100.0, 101.0], [104.0, 101.0]]
    assert np.all(vel == 0.0)


def test_head_on_impulse_same_cell():
    # v0 = +48, v1 = -48 along x, same cell.  Formulas: src/core.zig:354-391.
    pos, vel, out = run1([101.0, 101.0, 103.0, 101.0], [48.0, 0.0, -48.0, 0.0], collisions=True, capture=0)
    x0 = 101.0 + 48.0 * DT; x1 = 103.0 + (-48.0) * DT
    # first entry [i=1, j=0]
    dist = ((x0 - x1) * (x0 - x1) + 0.0) ** 0.5
    nx = (x0 - x1) / dist                       # -1.0
    van = (48.0 - (-48.0)) * nx + 0.0 * 0.0     # rel_vel = vj - vi = v0 - v1
    k = -(1.0 + 0.9) * van / 2
    overlap = 2.0 * 2.0 - dist
    corr = overlap * 0.5
    x1b = x1 - corr * nx; x0b = x0 + corr * nx
    v1b = -48.0 - nx * k; v0b = 48.0 + nx * k
    # second entry [i=0, j=1] with the corrected positions: dist is now 4 (overlap <= 0, no correction; Q3 impulse still applied)
    dist2 = ((x1b - x0b) * (x1b - x0b) + 0.0) ** 0.5
    nx2 = (x1b - x0b) / dist2
    van2 = (v1b - v0b) * nx2 + 0.0
    k2 = -(1.0 + 0.9) * van2 / 2
    overlap2 = 4.0 - dist2
    if overlap2 > 0:
        c2 = overlap2 * 0.5
        x0b, x1b = x0b - c2 * nx2, x1b + c2 * nx2
    v0c = v0b - nx2 * k2; v1c = v1b + nx2 * k2
    assert out["pairs"].tolist() == [[1, 0], [0, 1]]
    assert pos[0, 0] == x0b and pos[1, 0] == x1b
    assert vel[0, 0] == v0c and vel[1, 0] == v1c
    assert pos[0, 1] == 101.0 and pos[1, 1] == 101.0


def test_gravity_mode_uses_penalty_kick_not_position_correction():
    # Q5: t_acc switches the response to vel -/+= (10000*n*overlap)*dt and leaves positions alone
    # (src/core.zig:381-388).  Resting particles in ADJACENT cells (listed once), gravity on.
    pos, vel, out = run1([102.0, 101.0, 105.0, 101.0], [0.0, 0.0, 0.0, 0.0], collisions=True, t_acc=True, capture=0)
    vy = 0.0 + 50.0 * DT
    y = 101.0 + vy * DT
    # centre = particle 0 (cell (25,25)), neighbour = particle 1 via offset (1,0) -> entry [1, 0]
    assert out["pairs"].tolist() == [[1, 0]]
    dist = ((102.0 - 105.0) * (102.0 - 105.0) + (y - y) * (y - y)) ** 0.5
    nx = (102.0 - 105.0) / dist
    van = (0.0 - 0.0) * nx + (vy - vy) * 0.0
    k = -(1.0 + 0.9) * van / 2
    overlap = 4.0 - dist
    pax = (10000.0 * nx) * overlap
    v1x = 0.0 - pax * DT; v0x = 0.0 + pax * DT
    v1x = v1x - nx * k; v0x = v0x + nx * k
    assert pos.tolist() == [[102.0, y], [105.0, y]]
    assert vel[0, 0] == v0x and vel[1, 0] == v1x
    assert vel[0, 1] == vy and vel[1, 1] == vy


def test_pair_list_order_over_the_half_stencil():
    # A(0) in cell (1,1); B(1) in (1,2) [offset (0,1)]; C(2) in (2,1) [(1,0)]; D(3) in (0,2) [(-1,1)].
    # Cells are visited column-major (x slow), offsets in the order of src/core.zig:58.
    pos = [5.0, 7.5,   5.0, 8.5,   8.5, 7.5,   3.5, 8.5]
    _, _, out = run1(pos, [0.0] * 8, collisions=True, capture=0)
    assert out["pairs"].tolist() == [[1, 3], [1, 0], [2, 0], [3, 0], [1, 2]]


def test_coincident_particles_are_skipped():
    # dist < 1e-10 returns before any update (src/core.zig:355)
    pos, vel, out = run1([50.0, 50.0, 50.0, 50.0], [0.0] * 4, collisions=True, capture=0)
    assert out["pairs"].tolist() == [[1, 0], [0, 1]]
    assert pos.tolist() == [[50.0, 50.0], [50.0, 50.0]] and np.all(vel == 0.0)


def test_touching_is_not_a_collision():
    # strict d^2 < (2r)^2 (src/core.zig:342)
    _, _, out = run1([50.0, 50.0, 54.0, 50.0], [0.0] * 4, collisions=True, capture=0)
    assert out["pairs_count"] == 0


def test_last_cell_end_is_particle_count():
    # the last cell's end index is positions.len (src/core.zig:168)
    _, _, out = run1([717.0, 637.0, 718.0, 638.0], [0.0] * 4, collisions=True, capture=0)
    assert out["pairs"].tolist() == [[1, 0], [0, 1]]


def test_cell_population_stat_and_overflow_flag():
    pos = np.tile(np.array([[10.0, 10.0]]), (300, 1)) + np.linspace(0, 1e-3, 300)[:, None]
    _, _, out = run1(pos, np.zeros_like(pos), collisions=True)
    assert out["stats"]["max_cell_count"] == 300 and out["stats"]["cell_overflow"] == 1


def test_f32_oracle_rounds_every_operation_to_f32():
    pos, vel, _ = run1([100.0, 200.0], [48.5, -96.25], collisions=False, dtype=np.float32)
    dt = np.float32(np.float32(1) / np.float32(60)) / np.float32(8)
    assert pos[0, 0] == np.float32(100.0) + np.float32(48.5) * dt
    assert pos[0, 1] == np.float32(200.0) + np.float32(-96.25) * dt
