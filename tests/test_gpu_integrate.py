"""GPU parity, integrate + wall bounce (src/core.zig:248-261): libparticles_cuda through the C ABI
against the oracle on identical seeded inputs.  Bar: bit-exact positions and velocities, f64 and f32
(BASELINE.json configs[0], [1] and, through a sampled-particle property, [2])."""
import ctypes

import numpy as np
import pytest

from _util import DTYPES, assert_bit_exact, run_gpu, run_oracle
from particle_simulator_b200 import Simulation, capi, scenes

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("precision", ["f64", "f32"])
@pytest.mark.parametrize("t_acc", [False, True])
def test_default_scene_1000_substeps_bit_exact(precision, t_acc):
    scene = scenes.default_scene()
    pos0, vel0 = scenes.spawn(scene, seed=1, dtype=DTYPES[precision])
    want_p, want_v, _ = run_oracle(scene, pos0, vel0, 1000, collisions=False, t_acc=t_acc)
    got_p, got_v, _, stats = run_gpu(scene, pos0, vel0, 1000, precision=precision, collisions=False, t_acc=t_acc)
    assert stats["substeps_done"] == 1000
    assert_bit_exact(got_p, want_p, "positions")
    assert_bit_exact(got_v, want_v, "velocities")


@pytest.mark.parametrize("precision,t_acc", [("f64", False), ("f32", False), ("f32", True)])
def test_1m_particles_1000_substeps_bit_exact(precision, t_acc):
    scene = scenes.scaled_scene(1 << 20)
    pos0, vel0 = scenes.spawn(scene, seed=1, dtype=DTYPES[precision])
    want_p, want_v, _ = run_oracle(scene, pos0, vel0, 1000, collisions=False, t_acc=t_acc)
    got_p, got_v, _, _ = run_gpu(scene, pos0, vel0, 1000, precision=precision, collisions=False, t_acc=t_acc)
    assert_bit_exact(got_p, want_p, "positions")
    assert_bit_exact(got_v, want_v, "velocities")


@pytest.mark.parametrize("precision", ["f64", "f32"])
@pytest.mark.parametrize("n", [0, 1, 2, 3, 5, 255, 1023, 4097])
def test_ragged_counts(precision, n):
    scene = scenes.Scene(count=n, width=90.0, height=70.0)
    pos0, vel0 = scenes.spawn(scene, seed=4, dtype=DTYPES[precision])
    want_p, want_v, _ = run_oracle(scene, pos0, vel0, 50, collisions=False, t_acc=True)
    got_p, got_v, _, _ = run_gpu(scene, pos0, vel0, 50, precision=precision, collisions=False, t_acc=True)
    assert_bit_exact(got_p, want_p, "positions")
    assert_bit_exact(got_v, want_v, "velocities")


@pytest.mark.parametrize("precision", ["f64", "f32"])
@pytest.mark.parametrize("variant", [1, 2, 3, 4, 5, 6])
def test_every_kernel_variant_is_bit_identical(precision, variant):
    scene = scenes.Scene(count=100_003, width=900.0, height=700.0)
    pos0, vel0 = scenes.spawn(scene, seed=2, dtype=DTYPES[precision])
    want_p, want_v, _ = run_oracle(scene, pos0, vel0, 64, collisions=False, t_acc=True)
    with Simulation(scene, precision=precision, collisions=False) as sim:
        sim.debug_set_integrate_variant(variant)
        sim.upload(pos0, vel0)
        sim.set_acc(True)
        sim.step_substeps(64)
        got_p, got_v = sim.download(dtype=pos0.dtype)
    assert_bit_exact(got_p, want_p, "positions")
    assert_bit_exact(got_v, want_v, "velocities")


@pytest.mark.parametrize("precision", ["f64", "f32"])
@pytest.mark.parametrize("fuse", [2, 8, 5])
def test_register_fused_substeps_are_bit_identical(precision, fuse):
    scene = scenes.Scene(count=50_001, width=500.0, height=400.0)
    pos0, vel0 = scenes.spawn(scene, seed=3, dtype=DTYPES[precision])
    want_p, want_v, _ = run_oracle(scene, pos0, vel0, 83, collisions=False)
    got_p, got_v, _, stats = run_gpu(scene, pos0, vel0, 83, precision=precision, collisions=False, fuse_substeps=fuse)
    assert stats["substeps_done"] == 83
    assert_bit_exact(got_p, want_p, "positions")
    assert_bit_exact(got_v, want_v, "velocities")


def test_step_counts_frames_of_sub_steps():
    scene = scenes.default_scene().with_(count=1000)
    pos0, vel0 = scenes.spawn(scene, seed=1)
    want_p, want_v, _ = run_oracle(scene, pos0, vel0, 3 * 8, collisions=False)
    with Simulation(scene, collisions=False) as sim:
        sim.upload(pos0, vel0)
        sim.step(3)                     # 3 x updatePhysics = 24 sub-steps
        got_p, got_v = sim.download()
        assert sim.stats()["substeps_done"] == 24
    assert_bit_exact(got_p, want_p)
    assert_bit_exact(got_v, want_v)


def test_toggle_acc_between_frames():
    scene = scenes.default_scene().with_(count=2000)
    pos0, vel0 = scenes.spawn(scene, seed=6)
    p, v, _ = run_oracle(scene, pos0, vel0, 8, collisions=False, t_acc=False)
    p, v, _ = run_oracle(scene, p, v, 16, collisions=False, t_acc=True)
    p, v, _ = run_oracle(scene, p, v, 8, collisions=False, t_acc=False)
    with Simulation(scene, collisions=False) as sim:
        sim.upload(pos0, vel0)
        sim.step(1)
        sim.toggle_acc()
        assert sim.t_acc
        sim.step(2)
        acc = sim.download_acc()
        assert acc[0].tolist() == [0.0, 50.0]
        sim.toggle_acc()
        sim.step(1)
        got_p, got_v = sim.download()
    assert_bit_exact(got_p, p)
    assert_bit_exact(got_v, v)


def test_fresh_handle_is_zero_filled_and_roundtrips():
    scene = scenes.default_scene().with_(count=777)
    with Simulation(scene, precision="f32", collisions=False) as sim:
        p, v = sim.download()
        assert not p.any() and not v.any()              # Particles.init zero-fills, src/core.zig:27-29
        pos0, vel0 = scenes.spawn(scene, seed=9)
        sim.upload(pos0, vel0)                          # f64 host arrays into an f32 handle
        p64, v64 = sim.download()
        assert np.array_equal(p64, pos0.astype(np.float32).astype(np.float64))
        p32, v32 = sim.download(dtype=np.float32)       # the GUI's @floatCast read-back, src/main.zig:104-105
        assert np.array_equal(p32, pos0.astype(np.float32)) and np.array_equal(v32, vel0.astype(np.float32))
        sim.upload(None, vel0 * 0.5)                    # velocities only
        p2, v2 = sim.download(dtype=np.float32)
        assert np.array_equal(p2, p32) and np.array_equal(v2, (vel0 * 0.5).astype(np.float32))


def test_pinned_host_buffers_take_the_zero_copy_path():
    L = capi.load()
    scene = scenes.Scene(count=30_000, width=800.0, height=600.0)
    pos0, vel0 = scenes.spawn(scene, seed=8)
    nbytes = pos0.nbytes
    ptrs = [ctypes.c_void_p() for _ in range(4)]
    for p in ptrs:
        assert L.ps_host_alloc(ctypes.byref(p), nbytes) == capi.PS_OK
    try:
        arrs = [np.ctypeslib.as_array(ctypes.cast(p, ctypes.POINTER(ctypes.c_double)), shape=(scene.count, 2)) for p in ptrs]
        arrs[0][:] = pos0
        arrs[1][:] = vel0
        want_p, want_v, _ = run_oracle(scene, pos0, vel0, 16, collisions=False, t_acc=True)
        with Simulation(scene, collisions=False) as sim:
            sim.upload(arrs[0], arrs[1])
            sim.set_acc(True)
            sim.step(2)
            sim.download(arrs[2], arrs[3])
        assert_bit_exact(arrs[2].copy(), want_p)
        assert_bit_exact(arrs[3].copy(), want_v)
    finally:
        for p in ptrs:
            L.ps_host_free(p)


@pytest.mark.parametrize("precision", ["f32", "f64"])
def test_raw_kernel_on_torch_tensors(precision):
    """ps_integrate_bounce_* on caller-owned device memory (torch tensors, torch's current stream)."""
    import torch

    L = capi.load()
    scene = scenes.Scene(count=65_536 + 4, width=640.0, height=480.0)
    dt = DTYPES[precision]
    pos0, vel0 = scenes.spawn(scene, seed=5, dtype=dt)
    want_p, want_v, _ = run_oracle(scene, pos0, vel0, 24, collisions=False, t_acc=True)
    dev = torch.device("cuda:0")
    t = [torch.from_numpy(np.ascontiguousarray(a)).to(dev) for a in (pos0[:, 0], pos0[:, 1], vel0[:, 0], vel0[:, 1])]
    with Simulation(scene, precision=precision, collisions=False) as sim:
        cfg = sim.cfg
    fn = L.ps_integrate_bounce_f32 if precision == "f32" else L.ps_integrate_bounce_f64
    stream = torch.cuda.current_stream().cuda_stream
    rc = fn(ctypes.byref(cfg), 1, *[ctypes.c_void_p(x.data_ptr()) for x in t], scene.count, 24, ctypes.c_void_p(stream))
    assert rc == capi.PS_OK
    torch.cuda.synchronize()
    got_p = np.stack([t[0].cpu().numpy(), t[1].cpu().numpy()], axis=1)
    got_v = np.stack([t[2].cpu().numpy(), t[3].cpu().numpy()], axis=1)
    assert_bit_exact(got_p, want_p)
    assert_bit_exact(got_v, want_v)


def test_64m_particles_sampled_against_the_oracle():
    """BASELINE.json configs[2] at full size.  Particles are independent in this mode, so a random
    sample of them, fed to the oracle alone, must come out bit-identical to the full GPU run."""
    n = 64 << 20
    scene = scenes.scaled_scene(n)
    rng = np.random.default_rng(0)
    sample = np.sort(rng.choice(n, size=200_000, replace=False))
    substeps = 40
    with Simulation(scene, precision="f32", collisions=False) as sim:
        chunk = 8 << 20
        pos0 = np.empty((n, 2), dtype=np.float32)
        vel0 = np.empty((n, 2), dtype=np.float32)
        for s in range(0, n, chunk):
            p, v = scenes.spawn(scene, seed=1, start=s, count=chunk, dtype=np.float32)
            pos0[s:s + chunk] = p
            vel0[s:s + chunk] = v
        sim.upload(pos0, vel0)
        sim.set_acc(True)
        sim.step_substeps(substeps)
        got_p, got_v = sim.download(dtype=np.float32)
    sub = scene.with_(count=sample.size)
    want_p, want_v, _ = run_oracle(sub, pos0[sample], vel0[sample], substeps, collisions=False, t_acc=True)
    assert_bit_exact(got_p[sample], want_p, "sampled positions")
    assert_bit_exact(got_v[sample], want_v, "sampled velocities")
    # clamp invariant over the whole set (src/core.zig:316)
    assert got_p[:, 0].min() >= 2.0 and got_p[:, 0].max() <= np.float32(scene.width) - np.float32(2.0)
    assert got_p[:, 1].min() >= 2.0 and got_p[:, 1].max() <= np.float32(scene.height) - np.float32(2.0)


@pytest.mark.parametrize("precision", ["f64", "f32"])
def test_device_spawner_equals_host_generator(precision):
    """ps_spawn (Spawner.spawnRandom on the device) reproduces scenes.spawn bit for bit."""
    scene = scenes.Scene(count=200_003, width=3000.0, height=1800.0)
    want_p, want_v = scenes.spawn(scene, seed=11, dtype=DTYPES[precision])
    with Simulation(scene, precision=precision, collisions=False) as sim:
        sim.spawn(11)
        got_p, got_v = sim.download(dtype=DTYPES[precision])
    assert_bit_exact(got_p, want_p, "positions")
    assert_bit_exact(got_v, want_v, "velocities")


@pytest.mark.parametrize("precision", ["f32", "f64"])
def test_frame_host_streamed_equals_upload_step_download(precision):
    """ps_frame_host: updatePhysics on host slices.  With pinned buffers and an integration-only scene the particles are
    streamed in chunks over several streams; the result must be what upload + step + download gives (and the oracle)."""
    L = capi.load()
    n = (9 << 20) + 13                      # > 4 chunks, ragged tail
    scene = scenes.scaled_scene(n)
    ptrs = [ctypes.c_void_p() for _ in range(4)]
    for p in ptrs:
        assert L.ps_host_alloc(ctypes.byref(p), n * 16) == capi.PS_OK
    try:
        arrs = [np.ctypeslib.as_array(ctypes.cast(p, ctypes.POINTER(ctypes.c_double)), shape=(n, 2)) for p in ptrs]
        pos0, vel0 = scenes.spawn(scene, seed=3)
        arrs[0][:] = pos0
        arrs[1][:] = vel0
        with Simulation(scene, precision=precision, collisions=False) as sim:
            sim.set_acc(True)
            sim.frame_host(arrs[0], arrs[1], arrs[2], arrs[3], frames=2)
            streamed_p, streamed_v = arrs[2].copy(), arrs[3].copy()
            dev_p, dev_v = sim.download()                      # the device state must agree with what was written back
            assert sim.stats()["substeps_done"] == 16
        with Simulation(scene, precision=precision, collisions=False) as sim:
            sim.set_acc(True)
            sim.upload(pos0, vel0)
            sim.step(2)
            plain_p, plain_v = sim.download()
        assert_bit_exact(streamed_p, plain_p, "positions")
        assert_bit_exact(streamed_v, plain_v, "velocities")
        assert_bit_exact(dev_p, plain_p, "device positions")
        assert_bit_exact(dev_v, plain_v, "device velocities")
        rng = np.random.default_rng(1)
        sample = np.sort(rng.choice(n, size=50_000, replace=False))
        dt = DTYPES[precision]
        want_p, want_v, _ = run_oracle(scene.with_(count=sample.size), pos0[sample].astype(dt), vel0[sample].astype(dt), 16,
                                       collisions=False, t_acc=True)
        assert_bit_exact(streamed_p[sample].astype(dt), want_p, "sampled positions vs oracle")
        assert_bit_exact(streamed_v[sample].astype(dt), want_v, "sampled velocities vs oracle")
    finally:
        for p in ptrs:
            L.ps_host_free(p)


def test_frame_host_plain_route_with_collisions_and_pageable_buffers():
    scene = scenes.default_scene()
    pos0, vel0 = scenes.spawn(scene, seed=2)
    want_p, want_v, _ = run_oracle(scene, pos0, vel0, 24, collisions=True)
    out_p, out_v = np.empty_like(pos0), np.empty_like(vel0)
    with Simulation(scene, collisions=True) as sim:
        sim.frame_host(pos0, vel0, out_p, out_v, frames=3)
    assert_bit_exact(out_p, want_p)
    assert_bit_exact(out_v, want_v)
