"""oracle/core_ref.py — second, independent restatement of the reference step in pure Python.

TEST INFRASTRUCTURE ONLY.  PARITY UNPINNED (no Zig toolchain, the reference has no tests or
golden vectors): this file exists so that the C++ oracle (core_ref.hpp) is checked against
something written separately from the same source text, /root/reference/src/core.zig.
Python floats are IEEE f64 and every operation below is a single correctly rounded
operation, which is what Zig's strict float mode gives.  With ``real=numpy.float32`` every
operand is a float32 scalar and numpy rounds each operation to f32.

Pure-Python loops: small cases only (a few thousand particle-substeps).
"""
from __future__ import annotations

import math

import numpy as np

NEIGHBOR_OFFSETS = ((0, 0), (0, 1), (1, 0), (1, 1), (-1, 1))  # src/core.zig:58


class Particles:
    """src/core.zig:13-39 — positions/velocities/accelerations as lists of [x, y]."""

    def __init__(self, pos, vel, radius, real=float):
        self.real = real
        self.positions = [[real(p[0]), real(p[1])] for p in pos]
        self.velocities = [[real(v[0]), real(v[1])] for v in vel]
        self.accelerations = [[real(0.0), real(0.0)] for _ in pos]  # :29
        self.radius = real(radius)


class SpatialHashGrid:
    """src/core.zig:47-214."""

    def __init__(self, width, height, grid_size, count, real=float):
        self.real = real
        w, h, gs = int(width), int(height), int(grid_size)  # :61-63 (truncation)
        self.cols = w // gs  # :66 @divFloor
        self.rows = h // gs  # :67
        self.grid_size = real(grid_size)
        ncell = self.rows * self.cols
        self.grid_array = [0] * ncell
        self.cell_counters = [0] * ncell
        self.start_array = [0] * ncell
        self.sorted_array = [0] * count
        self.collision_pairs = []

    def clear(self):  # :95-100
        ncell = self.rows * self.cols
        self.grid_array = [0] * ncell
        self.cell_counters = [0] * ncell
        self.start_array = [0] * ncell
        self.sorted_array = [0] * len(self.sorted_array)

    def compute_coordinate(self, x, y):  # :103-115
        i = int(math.floor(x / self.grid_size))
        j = int(math.floor(y / self.grid_size))
        return min(i, self.cols - 1), min(j, self.rows - 1)

    def compute_grid_index(self, gx, gy):  # :119-121
        return gx * self.rows + gy

    def compute_grid_array(self, positions):  # :124-134
        for p in positions:
            gx, gy = self.compute_coordinate(p[0], p[1])
            self.grid_array[self.compute_grid_index(gx, gy)] += 1

    def compute_start_array(self):  # :137-142
        self.start_array[0] = 0
        for i in range(1, len(self.grid_array)):
            self.start_array[i] = self.start_array[i - 1] + self.grid_array[i - 1]

    def compute_sorted_array(self, positions):  # :145-159
        for i, p in enumerate(positions):
            gx, gy = self.compute_coordinate(p[0], p[1])
            c = self.compute_grid_index(gx, gy)
            self.sorted_array[self.start_array[c] + self.cell_counters[c]] = i
            self.cell_counters[c] += 1

    def detect_collisions(self, particles, dt, t_acc, particle_damping):  # :161-214
        self.collision_pairs = []
        ncell = self.cols * self.rows
        npart = len(particles.positions)
        for x in range(self.cols):
            for y in range(self.rows):
                c = self.compute_grid_index(x, y)
                s = self.start_array[c]
                e = self.start_array[c + 1] if c + 1 < ncell else npart
                if s >= e:
                    continue
                for j in range(s, e):
                    pj = self.sorted_array[j]
                    for dx, dy in NEIGHBOR_OFFSETS:
                        nx, ny = x + dx, y + dy
                        if nx < 0 or nx >= self.cols or ny < 0 or ny >= self.rows:
                            continue
                        nc = self.compute_grid_index(nx, ny)
                        s2 = self.start_array[nc]
                        e2 = self.start_array[nc + 1] if nc + 1 < ncell else npart
                        for i in range(s2, e2):
                            pi = self.sorted_array[i]
                            if pi == pj:
                                continue
                            if check_collision(particles.positions[pi], particles.positions[pj], particles.radius):
                                self.collision_pairs.append((pi, pj))
        for pi, pj in self.collision_pairs:  # :211-213
            resolve_collision(particles, pi, pj, dt, t_acc, particle_damping)


def distance_squared(xi, yi, xj, yj):  # :333-338
    dx_sq = (xj - xi) * (xj - xi)
    dy_sq = (yj - yi) * (yj - yi)
    return dx_sq + dy_sq


def check_collision(pos_i, pos_j, radius):  # :339-346
    real = type(radius)
    r_2 = real(2.0) * radius
    return distance_squared(pos_i[0], pos_i[1], pos_j[0], pos_j[1]) < r_2 * r_2


def _sqrt(x):
    return np.sqrt(x) if isinstance(x, np.floating) else math.sqrt(x)


def resolve_collision(particles, i, j, dt, t_acc, particle_damping):  # :348-392
    real = particles.real
    P, V = particles.positions, particles.velocities
    pi, pj = list(P[i]), list(P[j])
    vi, vj = list(V[i]), list(V[j])
    dist = _sqrt(distance_squared(pi[0], pi[1], pj[0], pj[1]))
    if dist < real(0.0000000001):
        return
    normal = [(pj[0] - pi[0]) / dist, (pj[1] - pi[1]) / dist]
    rel = [vj[0] - vi[0], vj[1] - vi[1]]
    van = rel[0] * normal[0] + rel[1] * normal[1]
    k = -(real(1.0) + particle_damping) * van / real(2)
    min_dis = real(2.0) * particles.radius
    overlap = min_dis - dist
    correction = overlap * real(0.5)
    if overlap > 0:
        if not t_acc:
            P[i] = [P[i][0] - correction * normal[0], P[i][1] - correction * normal[1]]
            P[j] = [P[j][0] + correction * normal[0], P[j][1] + correction * normal[1]]
        else:
            pen = [real(10000.0) * normal[0] * overlap, real(10000.0) * normal[1] * overlap]
            V[i] = [V[i][0] - pen[0] * dt, V[i][1] - pen[1] * dt]
            V[j] = [V[j][0] + pen[0] * dt, V[j][1] + pen[1] * dt]
    V[i] = [V[i][0] - normal[0] * k, V[i][1] - normal[1] * k]
    V[j] = [V[j][0] + normal[0] * k, V[j][1] + normal[1] * k]


class Solver:
    """src/core.zig:217-318."""

    def __init__(self, target_fps, sub_steps, g, boundary_damping, particle_damping, real=float):
        self.real = real
        self.delta_time = real(1) / real(target_fps)  # :227
        self.dt = self.delta_time / real(sub_steps)  # :228
        self.sub_steps = sub_steps
        self.g = real(g)
        self.boundary_damping = real(boundary_damping)
        self.particle_damping = real(particle_damping)
        self.t_acc = False  # :220

    def toggle_acc(self):  # :239-241
        self.t_acc = not self.t_acc

    def _zero_near_zero(self, x):  # :242-244
        return self.real(0) if abs(x) < self.real(1e-6) else x

    def substep(self, particles, width, height, shg, collisions=True):  # body of the loop at :247
        dt = self.dt
        for pos, vel, acc in zip(particles.positions, particles.velocities, particles.accelerations):
            if self.t_acc:
                acc[1] = self.g
            else:
                acc[0] = self.real(0.0)
                acc[1] = self.real(0.0)
            acc[0] = self._zero_near_zero(acc[0])
            acc[1] = self._zero_near_zero(acc[1])
            vel[0] = vel[0] + acc[0] * dt
            vel[1] = vel[1] + acc[1] * dt
            pos[0] = pos[0] + vel[0] * dt
            pos[1] = pos[1] + vel[1] * dt
        self._apply_boundary(particles, self.real(width), self.real(height))
        if not collisions:
            return
        shg.clear()
        shg.compute_grid_array(particles.positions)
        shg.compute_start_array()
        shg.compute_sorted_array(particles.positions)
        shg.detect_collisions(particles, self.dt, self.t_acc, self.particle_damping)

    def update_physics(self, particles, width, height, shg, collisions=True):  # :245-268
        for _ in range(self.sub_steps):
            self.substep(particles, width, height, shg, collisions)

    def _apply_boundary(self, particles, width, height):  # :290-318
        r = particles.radius
        lo = (r, r)
        hi = (width - r, height - r)
        neg = -self.boundary_damping
        for pos, vel in zip(particles.positions, particles.velocities):
            for c in (0, 1):
                m = (pos[c] < lo[c]) or (pos[c] > hi[c])
                if m:
                    vel[c] = vel[c] * neg
                pos[c] = min(max(pos[c], lo[c]), hi[c])

    def apply_boundary_scalar(self, particles, width, height):
        """The unused scalar variant, src/core.zig:269-288 — kept as a cross-check of _applyBoundary."""
        r = particles.radius
        for pos, vel in zip(particles.positions, particles.velocities):
            if pos[0] < r:
                pos[0] = r
                vel[0] *= -self.boundary_damping
            if pos[0] > width - r:
                pos[0] = width - r
                vel[0] *= -self.boundary_damping
            if pos[1] < r:
                pos[1] = r
                vel[1] *= -self.boundary_damping
            if pos[1] > height - r:
                pos[1] = height - r
                vel[1] *= -self.boundary_damping
