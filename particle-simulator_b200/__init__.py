"""particle-simulator_b200 — B200-native (sm_100a) particle step of J-A-G-A-N/Particle-Simulator.

The product is the C-ABI library ``libparticles_cuda.so`` (include/particles_cuda.h, csrc/).
This package holds its build recipe, the ctypes declarations, and the host-side mirror of
core.zig's API (``core.Particles``, ``core.Solver``, ``core.SpatialHashGrid``, ``core.Spawner``).
The directory name carries a hyphen; ``import particle_simulator_b200`` (the alias module at
the repository root) loads it.
"""
from . import build, capi, core, scenes  # noqa: F401
from .core import Particles, Simulation, SlabGroup, SlabSimulation, Solver, SpatialHashGrid, Spawner  # noqa: F401

__all__ = ["build", "capi", "core", "scenes", "Particles", "Simulation", "SlabGroup", "SlabSimulation", "Solver", "SpatialHashGrid", "Spawner"]
