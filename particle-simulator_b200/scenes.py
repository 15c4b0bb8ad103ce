"""Synthetic scenes: the main.zig constants and a reproducible stand-in for Spawner.spawnRandom.

The reference seeds its PRNG from the wall clock (src/main.zig:45), so it has no reproducible
initial state; parity always means "the same arrays fed to both sides".  The *distribution* is
the reference's (src/core.zig:403-408): x ~ U[r, W), y ~ U[r, H), vx, vy ~ U[min, max).  The
generator is counter-based (splitmix64 of seed and particle index) so every rank, the oracle
and the device spawner regenerate identical arrays without communication.
"""
from __future__ import annotations

from dataclasses import dataclass, replace

import numpy as np

# src/main.zig:25-29, 63-64, 72, 74
DEFAULT_COUNT = 13_000
DEFAULT_RADIUS = 2.0
DEFAULT_WIDTH = 720.0
DEFAULT_HEIGHT = 640.0
DEFAULT_FPS = 60.0
DEFAULT_SUB_STEPS = 8
DEFAULT_G = 50.0
DEFAULT_BOUNDARY_DAMPING = 0.1
DEFAULT_PARTICLE_DAMPING = 0.9
DEFAULT_VEL_MAX = 640.0  # @min(height, width), src/main.zig:67

# reference density: 13 000 particles over 180 x 160 cells of 4 px  
REFERENCE_DENSITY = 4.0 / 9.0  # particles per cell used for the scaled worlds (1M -> 1536^2 cells)


@dataclass(frozen=True)
class Scene:
    count: int = DEFAULT_COUNT
    radius: float = DEFAULT_RADIUS
    width: float = DEFAULT_WIDTH
    height: float = DEFAULT_HEIGHT
    grid_size: float = 2.0 * DEFAULT_RADIUS  # src/main.zig:74
    target_fps: float = DEFAULT_FPS
    sub_steps: int = DEFAULT_SUB_STEPS
    g: float = DEFAULT_G
    boundary_damping: float = DEFAULT_BOUNDARY_DAMPING
    particle_damping: float = DEFAULT_PARTICLE_DAMPING
    vel_min: float = -DEFAULT_VEL_MAX
    vel_max: float = DEFAULT_VEL_MAX

    def with_(self, **kw) -> "Scene":
        return replace(self, **kw)


def default_scene() -> Scene:
    """The scene main.zig builds (13 000 particles, 720x640)."""
    return Scene()


def scaled_scene(count: int) -> Scene:
    """A square world at the reference density (0.444 particles per 4-px cell), SURVEY §8(d):
    2^20 particles -> 6144^2 px, 16*2^20 -> 24576^2, 64*2^20 -> 49152^2, 256*2^20 -> 98304^2."""
    cells = count / REFERENCE_DENSITY
    side_cells = max(int(round(cells ** 0.5)), 1)
    side = float(side_cells * 4)
    return Scene(count=count, width=side, height=side)


_M64 = np.uint64(0xFFFFFFFFFFFFFFFF)


def splitmix64(z: np.ndarray) -> np.ndarray:
    """One splitmix64 output step on uint64 arrays (wrapping arithmetic)."""
    with np.errstate(over="ignore"):
        z = (z + np.uint64(0x9E3779B97F4A7C15)) & _M64
        z = ((z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)) & _M64
        z = ((z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)) & _M64
        return z ^ (z >> np.uint64(31))


def uniform01(seed: int, index: np.ndarray, stream: int) -> np.ndarray:
    """u = (splitmix64(splitmix64(seed) ^ (4*index + stream)) >> 11) * 2^-53  in [0, 1)."""
    base = splitmix64(np.array([seed], dtype=np.uint64))[0]
    ctr = (index.astype(np.uint64) * np.uint64(4) + np.uint64(stream)) ^ base
    return (splitmix64(ctr) >> np.uint64(11)).astype(np.float64) * (1.0 / 9007199254740992.0)


def spawn(scene: Scene, seed: int = 1, start: int = 0, count: int | None = None, dtype=np.float64):
    """Positions and velocities of particles [start, start+count) as interleaved (n, 2) arrays.

    lerp(a, b, t) = a + (b - a) * t, multiply and add rounded separately in f64, then cast.
    """
    n = scene.count - start if count is None else count
    idx = np.arange(start, start + n, dtype=np.uint64)
    r = scene.radius
    pos = np.empty((n, 2), dtype=np.float64)
    vel = np.empty((n, 2), dtype=np.float64)
    pos[:, 0] = r + (scene.width - r) * uniform01(seed, idx, 0)
    pos[:, 1] = r + (scene.height - r) * uniform01(seed, idx, 1)
    vel[:, 0] = scene.vel_min + (scene.vel_max - scene.vel_min) * uniform01(seed, idx, 2)
    vel[:, 1] = scene.vel_min + (scene.vel_max - scene.vel_min) * uniform01(seed, idx, 3)
    if dtype != np.float64:
        pos = pos.astype(dtype)
        vel = vel.astype(dtype)
    return pos, vel
