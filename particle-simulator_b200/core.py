"""Host-side mirror of src/core.zig's public API over the C ABI.

Same type and method names as the reference (``Particles``, ``SpatialHashGrid``, ``Solver``,
``Spawner``; ``updatePhysics``, ``toggle_acc``, ``spawnRandom``) so that a caller written
against core.zig — main.zig's init / updateAndRender (src/main.zig:35-76,95-114) — reads the
same.  All arithmetic happens in libparticles_cuda on the GPU; this file only moves arguments.

Differences a caller sees (documented in INTEGRATION.md):
  * ``Particles.positions`` / ``.velocities`` are host mirrors ((n, 2) float64, the layout of
    ``[]@Vector(2, f64)``).  ``Solver.updatePhysics`` refreshes ``positions`` after every call
    when ``read_back="frame"`` (what the GUI needs, src/main.zig:102-107); with
    ``read_back="never"`` (headless) call ``Particles.sync_host()`` when you want them.
  * writing into the host mirrors must be followed by ``Particles.mark_dirty()`` (Spawner does it).
  * ``SpatialHashGrid`` holds parameters only; its scratch arrays live on the device.
"""
from __future__ import annotations

import ctypes

import numpy as np

from . import capi, scenes


def _check(rc: int, handle=None) -> None:
    if rc != capi.PS_OK:
        L = capi.load()
        text = L.ps_last_error(handle) or L.ps_status_string(rc) or b""
        raise capi.PsError(rc, text.decode("utf-8", "replace"))


class Particles:
    """src/core.zig:13-45."""

    def __init__(self, particles_count: int, radius: float):
        self.positions = np.zeros((particles_count, 2), dtype=np.float64)      # :27
        self.velocities = np.zeros((particles_count, 2), dtype=np.float64)     # :28
        self.accelerations = np.zeros((particles_count, 2), dtype=np.float64)  # :29
        self.colors = np.zeros((particles_count, 4), dtype=np.uint8)           # sdl.Color r,g,b,a  :17
        self.radius = float(radius)
        self._dirty = True      # host mirrors newer than device state
        self._sim = None        # Simulation bound by Solver.updatePhysics

    @classmethod
    def init(cls, particles_count: int, radius: float) -> "Particles":
        return cls(particles_count, radius)

    def __len__(self) -> int:
        return self.positions.shape[0]

    def mark_dirty(self) -> None:
        self._dirty = True

    def sync_host(self, velocities: bool = True) -> None:
        """Copy the device state back into the host mirrors."""
        if self._sim is not None and not self._dirty:
            self._sim.download(self.positions, self.velocities if velocities else None)
            self._sim.download_acc(self.accelerations)

    def deinit(self) -> None:  # :41-44
        if self._sim is not None:
            self._sim.close()
            self._sim = None


class SpatialHashGrid:
    """src/core.zig:47-93 — parameters only; rows/cols derived as init does (:61-67)."""

    def __init__(self, width: float, height: float, grid_size: float, particles_count: int):
        self.width = float(width)
        self.height = float(height)
        self.grid_size = float(grid_size)
        self.particles_count = int(particles_count)
        w, h, gs = int(width), int(height), int(grid_size)
        if gs <= 0:
            raise ValueError("grid_size truncates to <= 0 (src/core.zig:63)")
        self.cols = w // gs
        self.rows = h // gs

    @classmethod
    def init(cls, width, height, grid_size, particles_count) -> "SpatialHashGrid":
        return cls(width, height, grid_size, particles_count)

    def deinit(self) -> None:
        pass


class Simulation:
    """Thin RAII wrapper of one ps_sim handle (one GPU)."""

    def __init__(self, scene: scenes.Scene, *, precision: str = "f64", collisions: bool = True, device: int = 0,
                 fuse_substeps: int = 0, contacts_per_particle: int = 0):
        self._L = capi.load()
        cfg = capi.PsConfig()
        cfg.struct_size = ctypes.sizeof(capi.PsConfig)
        cfg.device = device
        cfg.count = scene.count
        cfg.radius = scene.radius
        cfg.width = scene.width
        cfg.height = scene.height
        cfg.grid_size = scene.grid_size
        cfg.target_fps = scene.target_fps
        cfg.sub_steps = scene.sub_steps
        cfg.g = scene.g
        cfg.boundary_damping = scene.boundary_damping
        cfg.particle_damping = scene.particle_damping
        cfg.precision = capi.PS_F64 if precision == "f64" else capi.PS_F32
        cfg.collisions = 1 if collisions else 0
        cfg.fuse_substeps = fuse_substeps
        cfg.contacts_per_particle = contacts_per_particle
        self.cfg = cfg
        self.scene = scene
        self.precision = precision
        self.count = scene.count
        h = ctypes.c_void_p()
        rc = self._L.ps_create(ctypes.byref(cfg), ctypes.byref(h))
        if rc != capi.PS_OK:
            raise capi.PsError(rc, (self._L.ps_last_error(None) or b"").decode())
        self._h = h

    # -- lifetime
    def close(self) -> None:
        if getattr(self, "_h", None):
            self._L.ps_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    # -- state
    @staticmethod
    def _ptr(a):
        return None if a is None else ctypes.c_void_p(a.ctypes.data if isinstance(a, np.ndarray) else int(a))

    def upload(self, pos, vel) -> None:
        """pos/vel: C-contiguous (n, 2) float64 or float32 arrays (or None)."""
        for a in (pos, vel):
            if a is not None:
                assert a.flags.c_contiguous and a.shape == (self.count, 2), "expected a C-contiguous (count, 2) array"
        ref = pos if pos is not None else vel
        fn = self._L.ps_upload if ref.dtype == np.float64 else self._L.ps_upload_f32
        assert ref.dtype in (np.float64, np.float32)
        _check(fn(self._h, self._ptr(pos), self._ptr(vel)), self._h)

    def download(self, pos=None, vel=None, dtype=np.float64):
        """Fill (or allocate) interleaved (n, 2) arrays in original index order."""
        if pos is None and vel is None:
            pos = np.empty((self.count, 2), dtype=dtype)
            vel = np.empty((self.count, 2), dtype=dtype)
        ref = pos if pos is not None else vel
        fn = self._L.ps_download if ref.dtype == np.float64 else self._L.ps_download_f32
        _check(fn(self._h, self._ptr(pos), self._ptr(vel)), self._h)
        return pos, vel

    def spawn(self, seed: int, vel_min: float | None = None, vel_max: float | None = None) -> None:
        """Spawner.spawnRandom on the device (same values as scenes.spawn(scene, seed))."""
        vmin = self.scene.vel_min if vel_min is None else vel_min
        vmax = self.scene.vel_max if vel_max is None else vel_max
        _check(self._L.ps_spawn(self._h, seed, vmin, vmax), self._h)

    def export_rects(self) -> np.ndarray:
        """(n, 4) float32 {x - r, y - r, 2r, 2r}: the draw loop's SDL_FRect per particle (src/main.zig:102-107)."""
        out = np.empty((self.count, 4), dtype=np.float32)
        _check(self._L.ps_export_rects(self._h, self._ptr(out)), self._h)
        return out

    def set_colors(self, rgba: np.ndarray) -> None:
        assert rgba.dtype == np.uint8 and rgba.shape == (self.count, 4) and rgba.flags.c_contiguous
        _check(self._L.ps_set_colors(self._h, self._ptr(rgba)), self._h)

    def rainbow_colors(self) -> None:
        _check(self._L.ps_rainbow_colors(self._h), self._h)

    def get_colors(self) -> np.ndarray:
        out = np.empty((self.count, 4), dtype=np.uint8)
        _check(self._L.ps_get_colors(self._h, self._ptr(out)), self._h)
        return out

    def render(self, width: int, height: int) -> np.ndarray:
        """Headless frame (height, width, 4) uint8 in place of the SDL window."""
        out = np.empty((height, width, 4), dtype=np.uint8)
        _check(self._L.ps_render(self._h, width, height, self._ptr(out)), self._h)
        return out

    def download_acc(self, acc=None):
        if acc is None:
            acc = np.empty((self.count, 2), dtype=np.float64)
        _check(self._L.ps_download_acc(self._h, self._ptr(acc)), self._h)
        return acc

    # -- stepping
    def step(self, frames: int = 1) -> None:
        _check(self._L.ps_step(self._h, frames), self._h)

    def step_substeps(self, substeps: int) -> None:
        _check(self._L.ps_step_substeps(self._h, substeps), self._h)

    def frame_host(self, pos_in, vel_in, pos_out=None, vel_out=None, frames: int = 1) -> None:
        """updatePhysics on host slices: upload, `frames` frames, write back (ps_frame_host[_f32])."""
        fn = self._L.ps_frame_host if pos_in.dtype == np.float64 else self._L.ps_frame_host_f32
        _check(fn(self._h, self._ptr(pos_in), self._ptr(vel_in), self._ptr(pos_out), self._ptr(vel_out), frames), self._h)

    def toggle_acc(self) -> None:
        _check(self._L.ps_toggle_acc(self._h), self._h)

    def set_acc(self, on: bool) -> None:
        _check(self._L.ps_set_acc(self._h, int(on)), self._h)

    @property
    def t_acc(self) -> bool:
        return self._L.ps_get_acc(self._h) == 1

    def sync(self) -> None:
        _check(self._L.ps_sync(self._h), self._h)

    def set_stream(self, cuda_stream: int | None) -> None:
        _check(self._L.ps_set_stream(self._h, ctypes.c_void_p(cuda_stream or 0)), self._h)

    # -- inspection
    def grid_shape(self):
        r, c = ctypes.c_uint64(), ctypes.c_uint64()
        _check(self._L.ps_grid_shape(self._h, ctypes.byref(r), ctypes.byref(c)), self._h)
        return int(r.value), int(c.value)

    @property
    def dt(self) -> float:
        return float(self._L.ps_dt(self._h))

    def capture_pairs(self, on: bool = True) -> None:
        _check(self._L.ps_capture_pairs(self._h, int(on)), self._h)

    def contact_pairs(self) -> np.ndarray:
        """Ordered (m, 2) uint64 array {particle_i, particle_j} of the last sub-step."""
        cnt = ctypes.c_uint64()
        _check(self._L.ps_contact_pairs(self._h, None, 0, ctypes.byref(cnt)), self._h)
        out = np.empty((int(cnt.value), 2), dtype=np.uint64)
        if cnt.value:
            _check(self._L.ps_contact_pairs(self._h, self._ptr(out), cnt.value, ctypes.byref(cnt)), self._h)
        return out

    def stats(self) -> dict:
        s = capi.PsStats()
        _check(self._L.ps_get_stats(self._h, ctypes.byref(s)), self._h)
        return {k: int(getattr(s, k)) for k, _ in capi.PsStats._fields_ if k != "reserved"}

    def device_state(self):
        """Raw device pointers (px, py, vx, vy) of the SoA state."""
        p = [ctypes.c_void_p() for _ in range(4)]
        _check(self._L.ps_device_state(self._h, *[ctypes.byref(x) for x in p]), self._h)
        return tuple(int(x.value or 0) for x in p)

    def debug_set_integrate_variant(self, variant: int) -> None:
        _check(self._L.ps_debug_set_integrate_variant(self._h, variant), self._h)

    def debug_set_transfer_mode(self, mode: int) -> None:
        """0 = kernels read/write pinned host buffers directly; 1 = always DMA through device staging."""
        _check(self._L.ps_debug_set_transfer_mode(self._h, mode), self._h)


class Solver:
    """src/core.zig:217-268."""

    def __init__(self, target_fps: float, sub_steps: int, g: float, boundary_damping: float, particle_damping: float, *,
                 precision: str = "f64", collisions: bool = True, device: int = 0, read_back: str = "frame"):
        self.delta_time = 1.0 / target_fps                 # :227
        self.dt = self.delta_time / float(sub_steps)       # :228
        self.target_fps = float(target_fps)
        self.sub_steps = int(sub_steps)
        self.g = float(g)
        self.boundary_damping = float(boundary_damping)
        self.particle_damping = float(particle_damping)
        self.t_acc = False                                 # :220
        self.precision = precision
        self.collisions = collisions
        self.device = device
        self.read_back = read_back

    @classmethod
    def init(cls, target_fps, sub_steps, g, boundary_damping, particle_damping, **kw) -> "Solver":
        return cls(target_fps, sub_steps, g, boundary_damping, particle_damping, **kw)

    def toggle_acc(self) -> None:                          # :239-241
        self.t_acc = not self.t_acc

    def _bind(self, particles: Particles, width: float, height: float, shg: SpatialHashGrid) -> Simulation:
        scene = scenes.Scene(count=len(particles), radius=particles.radius, width=float(width), height=float(height),
                             grid_size=shg.grid_size, target_fps=self.target_fps, sub_steps=self.sub_steps, g=self.g,
                             boundary_damping=self.boundary_damping, particle_damping=self.particle_damping)
        sim = particles._sim
        if sim is None or sim.scene != scene or sim.precision != self.precision:
            if sim is not None:
                particles.sync_host()
                sim.close()
                particles._dirty = True
            sim = Simulation(scene, precision=self.precision, collisions=self.collisions, device=self.device)
            particles._sim = sim
        return sim

    def updatePhysics(self, particles: Particles, width: float, height: float, shg: SpatialHashGrid) -> None:
        """src/core.zig:245-268: sub_steps x (integrate, _applyBoundary, grid build, detect, resolve)."""
        sim = self._bind(particles, width, height, shg)
        if particles._dirty:
            sim.upload(particles.positions, particles.velocities)
            particles._dirty = False
        sim.set_acc(self.t_acc)
        sim.step(1)
        if self.read_back == "frame":
            sim.download(particles.positions, None)

    update_physics = updatePhysics


def hsv_to_rgba(h: np.ndarray, s: float, v: float, a: int) -> np.ndarray:
    """colors.zig hsvToSdlColor (src/colors.zig:6-47) in float32, vectorised over h."""
    h = h.astype(np.float32)
    c = np.float32(v) * np.float32(s)
    x = c * (np.float32(1.0) - np.abs(np.mod(h / np.float32(60.0), np.float32(2.0)) - np.float32(1.0)))
    m = np.float32(v) - c
    z = np.zeros_like(h)
    cc = np.full_like(h, c)
    sector = np.minimum((h // np.float32(60.0)).astype(np.int64), 5)
    r = np.choose(sector, [cc, x, z, z, x, cc])
    g = np.choose(sector, [x, cc, cc, x, z, z])
    b = np.choose(sector, [z, z, x, cc, cc, x])
    out = np.empty(h.shape + (4,), dtype=np.uint8)
    out[..., 0] = ((r + m) * np.float32(255.0)).astype(np.uint8)
    out[..., 1] = ((g + m) * np.float32(255.0)).astype(np.uint8)
    out[..., 2] = ((b + m) * np.float32(255.0)).astype(np.uint8)
    out[..., 3] = a
    return out


class Spawner:
    """src/core.zig:394-418.  `rng` is an integer seed: the reference's std.Random stream is
    not reproduced (its seed is the wall clock, src/main.zig:45); the distribution is."""

    @staticmethod
    def spawnRandom(par: Particles, rng: int, width: float, height: float, vmin: float, vmax: float) -> None:
        scene = scenes.Scene(count=len(par), radius=par.radius, width=float(width), height=float(height),
                             vel_min=float(vmin), vel_max=float(vmax))
        pos, vel = scenes.spawn(scene, seed=int(rng))
        par.positions[:] = pos
        par.velocities[:] = vel
        par.mark_dirty()
        Spawner.assignRainbowColors(par)

    @staticmethod
    def assignRainbowColors(par: Particles) -> None:      # :411-417
        n = len(par)
        if n == 0:
            return
        hue = np.arange(n, dtype=np.float32) / np.float32(n) * np.float32(360.0)
        par.colors[:] = hsv_to_rgba(hue, 1.0, 1.0, 255)

    spawn_random = spawnRandom


class SlabSimulation:
    """One rank of a collision scene decomposed into x-slabs across the GPUs of one box (ps_slab_*).

    `scene` is the WHOLE scene on every rank.  `nccl_id` is the 128-byte id from ``SlabSimulation.unique_id()``
    on rank 0, handed to all ranks by the host program (torch.distributed broadcast, a file, ...)."""

    def __init__(self, scene: scenes.Scene, rank: int, nranks: int, nccl_id: bytes | None = None, *, precision: str = "f64",
                 device: int = 0, contacts_per_particle: int = 0, capacity_percent: int = 0):
        self._L = capi.load()
        cfg = capi.PsConfig()
        cfg.struct_size = ctypes.sizeof(capi.PsConfig)
        cfg.device = device
        cfg.count = scene.count
        cfg.radius, cfg.width, cfg.height, cfg.grid_size = scene.radius, scene.width, scene.height, scene.grid_size
        cfg.target_fps, cfg.sub_steps, cfg.g = scene.target_fps, scene.sub_steps, scene.g
        cfg.boundary_damping, cfg.particle_damping = scene.boundary_damping, scene.particle_damping
        cfg.precision = capi.PS_F64 if precision == "f64" else capi.PS_F32
        cfg.collisions = 1
        cfg.contacts_per_particle = contacts_per_particle
        cfg.slab_capacity_percent = capacity_percent
        self.cfg, self.scene, self.rank, self.nranks = cfg, scene, rank, nranks
        h = ctypes.c_void_p()
        idbuf = ctypes.create_string_buffer(nccl_id, 128) if nccl_id is not None else None
        rc = self._L.ps_slab_create(ctypes.byref(cfg), rank, nranks, idbuf, ctypes.byref(h))
        if rc != capi.PS_OK:
            raise capi.PsError(rc, (self._L.ps_slab_last_error(None) or b"").decode())
        self._h = h

    @staticmethod
    def unique_id() -> bytes:
        L = capi.load()
        buf = ctypes.create_string_buffer(128)
        rc = L.ps_nccl_unique_id(buf)
        if rc != capi.PS_OK:
            raise capi.PsError(rc, (L.ps_slab_last_error(None) or b"").decode())
        return buf.raw

    def _check(self, rc):
        if rc != capi.PS_OK:
            raise capi.PsError(rc, (self._L.ps_slab_last_error(self._h) or self._L.ps_status_string(rc) or b"").decode())

    def close(self):
        if getattr(self, "_h", None):
            self._L.ps_slab_destroy(self._h)
            self._h = None

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def upload_global(self, pos: np.ndarray, vel: np.ndarray) -> None:
        assert pos.dtype == np.float64 and vel.dtype == np.float64 and pos.shape == (self.scene.count, 2) == vel.shape
        assert pos.flags.c_contiguous and vel.flags.c_contiguous
        self._check(self._L.ps_slab_upload_global(self._h, ctypes.c_void_p(pos.ctypes.data), ctypes.c_void_p(vel.ctypes.data)))

    def spawn(self, seed: int, vel_min: float | None = None, vel_max: float | None = None) -> None:
        vmin = self.scene.vel_min if vel_min is None else vel_min
        vmax = self.scene.vel_max if vel_max is None else vel_max
        self._check(self._L.ps_slab_spawn(self._h, seed, vmin, vmax))

    def step(self, frames: int = 1) -> None:
        self._check(self._L.ps_slab_step(self._h, frames))

    def step_substeps(self, substeps: int) -> None:
        self._check(self._L.ps_slab_step_substeps(self._h, substeps))

    def set_acc(self, on: bool) -> None:
        self._check(self._L.ps_slab_set_acc(self._h, int(on)))

    def sync(self) -> None:
        self._check(self._L.ps_slab_sync(self._h))

    def set_stream(self, cuda_stream: int | None) -> None:
        self._check(self._L.ps_slab_set_stream(self._h, ctypes.c_void_p(cuda_stream or 0)))

    TIMING_STAGES = ("bin", "M0", "sort+detect", "resolve A+check", "read-back", "M1", "phase B", "finish")

    def set_timing(self, on: bool) -> None:
        self._check(self._L.ps_slab_set_timing(self._h, int(on)))

    def timing(self) -> dict:
        """Microseconds per sub-step of every stage on this rank since set_timing(True) (first 8 sub-steps discarded)."""
        us = (ctypes.c_double * len(self.TIMING_STAGES))()
        n = ctypes.c_uint64()
        self._check(self._L.ps_slab_get_timing(self._h, us, len(self.TIMING_STAGES), ctypes.byref(n)))
        out = {k: float(us[i]) for i, k in enumerate(self.TIMING_STAGES)}
        out["substeps"] = int(n.value)
        return out

    def columns(self):
        a, b = ctypes.c_uint64(), ctypes.c_uint64()
        self._check(self._L.ps_slab_columns(self._h, ctypes.byref(a), ctypes.byref(b)))
        return int(a.value), int(b.value)

    def owned_count(self) -> int:
        n = ctypes.c_uint64()
        self._check(self._L.ps_slab_owned_count(self._h, ctypes.byref(n)))
        return int(n.value)

    def download_owned(self):
        """(ids uint64 (m,), pos float64 (m, 2), vel float64 (m, 2)) of the particles this rank owns."""
        m = self.owned_count()
        ids = np.empty(max(m, 1), dtype=np.uint64)
        pos = np.empty((max(m, 1), 2), dtype=np.float64)
        vel = np.empty((max(m, 1), 2), dtype=np.float64)
        cnt = ctypes.c_uint64()
        self._check(self._L.ps_slab_download_owned(self._h, ctypes.c_void_p(ids.ctypes.data), ctypes.c_void_p(pos.ctypes.data),
                                                   ctypes.c_void_p(vel.ctypes.data), max(m, 1), ctypes.byref(cnt)))
        m = int(cnt.value)
        return ids[:m], pos[:m], vel[:m]

    def stats(self) -> dict:
        s = capi.PsStats()
        self._check(self._L.ps_slab_get_stats(self._h, ctypes.byref(s)))
        return {k: int(getattr(s, k)) for k, _ in capi.PsStats._fields_ if k != "reserved"}


class SlabGroup:
    """All ranks of a slab-decomposed scene emulated in one process on one GPU (ps_slabgroup_*): the multi-rank
    protocol with device copies instead of NCCL.  For tests and for boxes with fewer GPUs than slabs."""

    def __init__(self, scene: scenes.Scene, nranks: int, *, precision: str = "f64", device: int = 0, contacts_per_particle: int = 0,
                 capacity_percent: int = 0):
        self._L = capi.load()
        cfg = capi.PsConfig()
        cfg.struct_size = ctypes.sizeof(capi.PsConfig)
        cfg.device = device
        cfg.count = scene.count
        cfg.radius, cfg.width, cfg.height, cfg.grid_size = scene.radius, scene.width, scene.height, scene.grid_size
        cfg.target_fps, cfg.sub_steps, cfg.g = scene.target_fps, scene.sub_steps, scene.g
        cfg.boundary_damping, cfg.particle_damping = scene.boundary_damping, scene.particle_damping
        cfg.precision = capi.PS_F64 if precision == "f64" else capi.PS_F32
        cfg.collisions = 1
        cfg.contacts_per_particle = contacts_per_particle
        cfg.slab_capacity_percent = capacity_percent
        self.scene, self.nranks = scene, nranks
        h = ctypes.c_void_p()
        rc = self._L.ps_slabgroup_create(ctypes.byref(cfg), nranks, ctypes.byref(h))
        if rc != capi.PS_OK:
            raise capi.PsError(rc, (self._L.ps_slabgroup_last_error(None) or b"").decode())
        self._h = h

    def _check(self, rc):
        if rc != capi.PS_OK:
            raise capi.PsError(rc, (self._L.ps_slabgroup_last_error(self._h) or self._L.ps_status_string(rc) or b"").decode())

    def close(self):
        if getattr(self, "_h", None):
            self._L.ps_slabgroup_destroy(self._h)
            self._h = None

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def upload_global(self, pos, vel):
        assert pos.dtype == np.float64 and pos.flags.c_contiguous and vel.flags.c_contiguous
        self._check(self._L.ps_slabgroup_upload_global(self._h, ctypes.c_void_p(pos.ctypes.data), ctypes.c_void_p(vel.ctypes.data)))

    def set_acc(self, on):
        self._check(self._L.ps_slabgroup_set_acc(self._h, int(on)))

    def step_substeps(self, n):
        self._check(self._L.ps_slabgroup_step_substeps(self._h, n))

    def download_owned(self, rank):
        n = ctypes.c_uint64()
        self._check(self._L.ps_slabgroup_owned_count(self._h, rank, ctypes.byref(n)))
        m = max(int(n.value), 1)
        ids = np.empty(m, dtype=np.uint64); pos = np.empty((m, 2)); vel = np.empty((m, 2))
        cnt = ctypes.c_uint64()
        self._check(self._L.ps_slabgroup_download_owned(self._h, rank, ctypes.c_void_p(ids.ctypes.data), ctypes.c_void_p(pos.ctypes.data),
                                                        ctypes.c_void_p(vel.ctypes.data), m, ctypes.byref(cnt)))
        m = int(cnt.value)
        return ids[:m], pos[:m], vel[:m]

    def stats(self, rank):
        s = capi.PsStats()
        self._check(self._L.ps_slabgroup_get_stats(self._h, rank, ctypes.byref(s)))
        return {k: int(getattr(s, k)) for k, _ in capi.PsStats._fields_ if k != "reserved"}
