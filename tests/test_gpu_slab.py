"""GPU parity of the slab-decomposed collision step (ps_slab_*, csrc/ps_slab.cuh) against the oracle.

One rank on one GPU exercises the slab kernels without a neighbour; two ranks (one process per GPU, NCCL
over NVLink) exercise migration, halo, blocked boundary particles and the M1 / M2 hand-over.  The two-rank
tests need two GPUs (gpurun --gpus 2) and skip themselves otherwise."""
import os

import numpy as np
import pytest

from _util import assert_bit_exact, run_oracle
from particle_simulator_b200 import SlabSimulation, scenes

pytestmark = pytest.mark.gpu


def gather(parts, n, dtype):
    pos = np.full((n, 2), np.nan, dtype=np.float64)
    vel = np.full((n, 2), np.nan, dtype=np.float64)
    seen = np.zeros(n, dtype=np.int64)
    for ids, p, v in parts:
        pos[ids.astype(np.int64)] = p
        vel[ids.astype(np.int64)] = v
        seen[ids.astype(np.int64)] += 1
    assert (seen == 1).all(), "every particle must be owned by exactly one rank"
    return pos.astype(dtype), vel.astype(dtype)


@pytest.mark.parametrize("precision,dtype", [("f64", np.float64), ("f32", np.float32)])
@pytest.mark.parametrize("t_acc", [False, True])
def test_single_rank_slab_equals_oracle(precision, dtype, t_acc):
    scene = scenes.default_scene()
    pos0, vel0 = scenes.spawn(scene, seed=1)
    want_p, want_v, out = run_oracle(scene, pos0.astype(dtype), vel0.astype(dtype), 200, collisions=True, t_acc=t_acc)
    with SlabSimulation(scene, 0, 1, None, precision=precision) as s:
        s.upload_global(pos0, vel0)
        s.set_acc(t_acc)
        s.step_substeps(200)
        s.sync()
        parts = [s.download_owned()]
        st = s.stats()
    got_p, got_v = gather(parts, scene.count, dtype)
    assert st["pairs_total"] == out["stats"]["pairs_total"]
    assert_bit_exact(got_p, want_p, "positions")
    assert_bit_exact(got_v, want_v, "velocities")


def test_device_spawner_matches_host_generator():
    scene = scenes.default_scene().with_(count=50_000, width=1200.0, height=900.0)
    pos0, vel0 = scenes.spawn(scene, seed=7)
    with SlabSimulation(scene, 0, 1, None, precision="f64") as s:
        s.spawn(7)
        ids, p, v = s.download_owned()
    got_p, got_v = gather([(ids, p, v)], scene.count, np.float64)
    assert_bit_exact(got_p, pos0)
    assert_bit_exact(got_v, vel0)


@pytest.mark.parametrize("nranks,precision,t_acc,seed", [(2, "f64", False, 1), (3, "f32", True, 2), (4, "f64", True, 3), (8, "f32", False, 4)])
def test_ranks_emulated_on_one_gpu_equal_oracle(nranks, precision, t_acc, seed):
    """The whole multi-rank protocol (migration, halo, blocked boundary particles, M1 / M2 hand-over) with all
    ranks in one process on one GPU (SlabGroup): same kernels as the NCCL path, device copies for messages."""
    from particle_simulator_b200 import SlabGroup

    dtype = np.float64 if precision == "f64" else np.float32
    scene = scenes.default_scene()
    pos0, vel0 = scenes.spawn(scene, seed=seed)
    want_p, want_v, out = run_oracle(scene, pos0.astype(dtype), vel0.astype(dtype), 300, collisions=True, t_acc=t_acc)
    with SlabGroup(scene, nranks, precision=precision, contacts_per_particle=32) as g:
        g.upload_global(pos0, vel0)
        g.set_acc(t_acc)
        g.step_substeps(300)
        parts = [g.download_owned(r) for r in range(nranks)]
        pairs = sum(g.stats(r)["pairs_total"] for r in range(nranks))
    got_p, got_v = gather(parts, scene.count, dtype)
    assert pairs == out["stats"]["pairs_total"]
    assert_bit_exact(got_p, want_p, "positions")
    assert_bit_exact(got_v, want_v, "velocities")


def test_emulated_ranks_dense_narrow_slabs():
    from particle_simulator_b200 import SlabGroup

    scene = scenes.Scene(count=3000, width=64.0, height=400.0)       # 16 columns, 4 ranks: 4 columns each, ~1.9 particles per cell
    pos0, vel0 = scenes.spawn(scene, seed=5)
    want_p, want_v, out = run_oracle(scene, pos0, vel0, 120, collisions=True, t_acc=False)
    with SlabGroup(scene, 4, precision="f64", contacts_per_particle=64) as g:
        g.upload_global(pos0, vel0)
        g.step_substeps(120)
        parts = [g.download_owned(r) for r in range(4)]
    got_p, got_v = gather(parts, scene.count, np.float64)
    assert_bit_exact(got_p, want_p, "positions")
    assert_bit_exact(got_v, want_v, "velocities")


def dense_band_scene(count=6000, seed=9):
    """Far from uniform: most particles in a band astride the boundary between two slabs (8 ranks of the default world:
    22-23 cell columns each, boundaries at x = 88, 180, 268, ...), falling under gravity."""
    scene = scenes.default_scene().with_(count=count)
    rng = np.random.default_rng(seed)
    pos = np.empty((count, 2))
    vel = rng.uniform(-200.0, 200.0, size=(count, 2))
    band = rng.random(count) < 0.8
    pos[:, 0] = np.where(band, np.clip(rng.normal(180.0, 25.0, count), scene.radius, scene.width - scene.radius),
                         rng.uniform(scene.radius, scene.width, count))
    pos[:, 1] = rng.uniform(scene.radius, scene.height, count)
    return scene, np.ascontiguousarray(pos), np.ascontiguousarray(vel)


@pytest.mark.parametrize("nranks,precision", [(8, "f64"), (4, "f32")])
def test_emulated_ranks_dense_band_under_gravity(nranks, precision):
    """Non-uniform load: a dense band on a slab boundary with gravity on (the penalty branch of resolve_collision, long
    dependency chains that run across the boundary, ranks holding several times the average share)."""
    from particle_simulator_b200 import SlabGroup

    dtype = np.float64 if precision == "f64" else np.float32
    scene, pos0, vel0 = dense_band_scene()
    want_p, want_v, out = run_oracle(scene, pos0.astype(dtype), vel0.astype(dtype), 480, collisions=True, t_acc=True)
    with SlabGroup(scene, nranks, precision=precision, contacts_per_particle=96) as g:
        g.upload_global(pos0, vel0)
        g.set_acc(True)
        g.step_substeps(480)
        parts = [g.download_owned(r) for r in range(nranks)]
        pairs = sum(g.stats(r)["pairs_total"] for r in range(nranks))
    got_p, got_v = gather(parts, scene.count, dtype)
    assert pairs == out["stats"]["pairs_total"]
    assert_bit_exact(got_p, want_p, "positions")
    assert_bit_exact(got_v, want_v, "velocities")
    counts = [len(ids) for ids, _, _ in parts]
    assert max(counts) > 3 * (scene.count // nranks) // 2, "the scene was meant to be unbalanced"


def test_slab_capacity_is_configurable():
    """Above 4 Mi particles a rank is sized for 1.25 x the average share (ps_config.slab_capacity_percent = 0).  A scene with
    nine tenths of its particles in one rank's slab is refused with PS_ERR_CAPACITY at that size, and runs, bit-identical to the
    oracle, once the caller asks for room."""
    from particle_simulator_b200 import SlabGroup, capi

    n = 5_000_000
    scene = scenes.scaled_scene(n)
    rng = np.random.default_rng(21)
    pos = np.empty((n, 2))
    left = rng.random(n) < 0.9
    pos[:, 0] = np.where(left, rng.uniform(scene.radius, scene.width * 0.45, n), rng.uniform(scene.width * 0.55, scene.width - scene.radius, n))
    pos[:, 1] = rng.uniform(scene.radius, scene.height - scene.radius, n)
    vel = rng.uniform(-640.0, 640.0, size=(n, 2))
    with pytest.raises(capi.PsError) as e:
        with SlabGroup(scene, 2, precision="f32") as g:
            g.upload_global(pos, vel)
    assert e.value.status == capi.PS_ERR_CAPACITY
    want_p, want_v, out = run_oracle(scene, pos.astype(np.float32), vel.astype(np.float32), 2, collisions=True)
    with SlabGroup(scene, 2, precision="f32", capacity_percent=200, contacts_per_particle=24) as g:
        g.upload_global(pos, vel)
        g.step_substeps(2)
        parts = [g.download_owned(r) for r in range(2)]
    got_p, got_v = gather(parts, n, np.float32)
    assert_bit_exact(got_p, want_p, "positions")
    assert_bit_exact(got_v, want_v, "velocities")
    assert len(parts[0][0]) > 4_000_000


def _two_rank_worker(rank, world, port, tmpdir, case):
    import torch
    import torch.distributed as dist

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("gloo", rank=rank, world_size=world)   # host-side channel for the NCCL id only
    idt = torch.zeros(128, dtype=torch.uint8)
    if rank == 0:
        idt = torch.frombuffer(bytearray(SlabSimulation.unique_id()), dtype=torch.uint8).clone()
    dist.broadcast(idt, 0)
    scene, seed, substeps, precision, t_acc = case[:5]
    if len(case) > 5 and case[5]:
        os.environ["PS_SLAB_FORCE_LATE"] = "1"      # every rank sends its boundary states only after its own phase B
    if isinstance(scene, str):
        scene, pos0, vel0 = dense_band_scene(seed=seed)
    else:
        pos0, vel0 = scenes.spawn(scene, seed=seed)
    with SlabSimulation(scene, rank, world, bytes(idt.numpy().tobytes()), precision=precision, device=rank, contacts_per_particle=96) as s:
        s.upload_global(pos0, vel0)
        s.set_acc(t_acc)
        s.step_substeps(substeps)
        s.sync()
        ids, p, v = s.download_owned()
        st = s.stats()
    np.savez(os.path.join(tmpdir, f"rank{rank}.npz"), ids=ids, pos=p, vel=v, pairs_total=st["pairs_total"])
    dist.barrier()
    dist.destroy_process_group()


CASES = {
    "default_f64": (scenes.default_scene(), 1, 400, "f64", False),
    "default_f32_gravity": (scenes.default_scene(), 2, 400, "f32", True),
    "dense_narrow": (scenes.Scene(count=3000, width=64.0, height=400.0), 3, 150, "f64", False),   # 16 columns: 8 per rank
    "one_million": (scenes.scaled_scene(1 << 20), 1, 24, "f32", False),
    "late_boundary_path": (scenes.default_scene(), 5, 120, "f64", True, True),
    "dense_band_gravity": ("dense_band", 9, 320, "f64", True),
}


@pytest.mark.parametrize("name", list(CASES))
@pytest.mark.parametrize("world", [2])
def test_two_ranks_equal_oracle(name, world, tmp_path):
    import torch
    import torch.multiprocessing as mp

    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    case = CASES[name]
    scene, seed, substeps, precision, t_acc = case[:5]
    dtype = np.float64 if precision == "f64" else np.float32
    port = 29600 + (os.getpid() % 1000)
    mp.spawn(_two_rank_worker, args=(world, port, str(tmp_path), case), nprocs=world, join=True)
    if isinstance(scene, str):
        scene, pos0, vel0 = dense_band_scene(seed=seed)
    else:
        pos0, vel0 = scenes.spawn(scene, seed=seed)
    want_p, want_v, out = run_oracle(scene, pos0.astype(dtype), vel0.astype(dtype), substeps, collisions=True, t_acc=t_acc)
    parts, pairs = [], 0
    for r in range(world):
        z = np.load(tmp_path / f"rank{r}.npz")
        parts.append((z["ids"], z["pos"], z["vel"]))
        pairs += int(z["pairs_total"])
    got_p, got_v = gather(parts, scene.count, dtype)
    assert pairs == out["stats"]["pairs_total"]
    assert_bit_exact(got_p, want_p, "positions")
    assert_bit_exact(got_v, want_v, "velocities")
