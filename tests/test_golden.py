"""Frozen oracle outputs (tests/golden/, written by tests/golden/make_golden.py).

CPU tier: the oracle still reproduces them bit-for-bit (guards the checker itself).
GPU tier: libparticles_cuda reproduces them through the C ABI without the oracle being run at all."""
import hashlib
import json
import os

import numpy as np
import pytest

import oracle
from particle_simulator_b200 import scenes

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
SMALL = ["small_collide", "small_collide_gravity", "small_integrate_gravity"]
DTYPES = {"f64": np.float64, "f32": np.float32}


def load(name, tag):
    z = np.load(os.path.join(GOLDEN, f"{name}_{tag}.npz"))
    meta = json.loads(str(z["meta"]))
    scene = scenes.Scene(count=meta["count"], width=meta["width"], height=meta["height"])
    return z, meta, scene


def digest(*arrays):
    h = hashlib.sha256()
    for a in arrays:
        h.update(np.ascontiguousarray(a).tobytes())
    return h.hexdigest()


@pytest.mark.parametrize("tag", ["f64", "f32"])
@pytest.mark.parametrize("name", SMALL)
def test_oracle_reproduces_golden(name, tag):
    z, meta, scene = load(name, tag)
    pos, vel = scenes.spawn(scene, seed=meta["seed"], dtype=DTYPES[tag])
    p = oracle.make_params(width=scene.width, height=scene.height)
    out = oracle.run(p, pos, vel, meta["substeps"], collisions=meta["collisions"], t_acc=meta["t_acc"],
                     capture_substep=meta["substeps"] - 1 if meta["collisions"] else -1, pairs_cap=16 * scene.count)
    assert np.array_equal(pos.view(np.uint8), z["pos"].view(np.uint8))
    assert np.array_equal(vel.view(np.uint8), z["vel"].view(np.uint8))
    if meta["collisions"]:
        assert np.array_equal(out["pairs"].astype(np.uint32), z["pairs"])
        assert out["stats"]["pairs_total"] == meta["pairs_total"]


@pytest.mark.parametrize("key", ["default_integrate_f64", "default_integrate_f32"])
def test_oracle_reproduces_default_scene_digest(key):
    d = json.load(open(os.path.join(GOLDEN, "default_scene_digests.json")))[key]
    dt = np.float64 if key.endswith("f64") else np.float32
    scene = scenes.default_scene()
    pos, vel = scenes.spawn(scene, seed=d["seed"], dtype=dt)
    oracle.run(oracle.make_params(), pos, vel, d["substeps"], collisions=d["collisions"], t_acc=d["t_acc"])
    assert digest(pos, vel) == d["sha256_pos_vel"]


@pytest.mark.gpu
@pytest.mark.parametrize("tag", ["f64", "f32"])
@pytest.mark.parametrize("name", SMALL)
def test_gpu_reproduces_golden(name, tag):
    from particle_simulator_b200 import Simulation

    z, meta, scene = load(name, tag)
    pos0, vel0 = scenes.spawn(scene, seed=meta["seed"], dtype=DTYPES[tag])
    with Simulation(scene, precision=tag, collisions=meta["collisions"]) as sim:
        sim.upload(pos0, vel0)
        sim.set_acc(meta["t_acc"])
        if meta["collisions"]:
            sim.capture_pairs(True)
        sim.step_substeps(meta["substeps"])
        pos, vel = sim.download(dtype=DTYPES[tag])
        if meta["collisions"]:
            assert np.array_equal(sim.contact_pairs().astype(np.uint32), z["pairs"])
            assert sim.stats()["pairs_total"] == meta["pairs_total"]
    assert np.array_equal(pos.view(np.uint8), z["pos"].view(np.uint8))
    assert np.array_equal(vel.view(np.uint8), z["vel"].view(np.uint8))


@pytest.mark.gpu
def test_gpu_reproduces_default_scene_digests():
    from particle_simulator_b200 import Simulation

    digests = json.load(open(os.path.join(GOLDEN, "default_scene_digests.json")))
    scene = scenes.default_scene()
    for key, d in digests.items():
        tag = key[-3:]
        pos0, vel0 = scenes.spawn(scene, seed=d["seed"], dtype=DTYPES[tag])
        with Simulation(scene, precision=tag, collisions=d["collisions"]) as sim:
            sim.upload(pos0, vel0)
            sim.set_acc(d["t_acc"])
            sim.step_substeps(d["substeps"])
            pos, vel = sim.download(dtype=DTYPES[tag])
            if d["collisions"]:
                assert sim.stats()["pairs_total"] == d["pairs_total"]
        assert digest(pos, vel) == d["sha256_pos_vel"], key
