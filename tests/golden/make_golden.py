#!/usr/bin/env python
"""Generate tests/golden/*.npz — frozen oracle outputs on seeded scenes.

The reference publishes no golden vectors and cannot be run here (Zig), so these are the ORACLE's outputs
(oracle/core_ref.hpp, f64 and f32), frozen after the oracle passed its hand-derived known-answer tests and agreed
bit-for-bit with the independent Python restatement.  They guard the oracle against regressions (CPU tier) and
give the GPU tier a fixture that does not depend on running the oracle.  Parity with the Zig binary stays unpinned.

    python tests/golden/make_golden.py          # rewrites the fixtures (run from the repository root)
"""
import hashlib
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

import oracle  # noqa: E402
from particle_simulator_b200 import scenes  # noqa: E402

SMALL = scenes.Scene(count=2000, width=280.0, height=260.0)   # ~0.44 particles per cell, like the default scene
CASES = [  # name, scene, seed, substeps, collisions, t_acc
    ("small_collide", SMALL, 11, 400, True, False),
    ("small_collide_gravity", SMALL, 12, 400, True, True),
    ("small_integrate_gravity", SMALL, 13, 1000, False, True),
]
DEFAULT_CASES = [  # full default scene: only digests are stored
    ("default_collide", scenes.default_scene(), 1, 1000, True, False),
    ("default_collide_gravity", scenes.default_scene(), 1, 1000, True, True),
    ("default_integrate", scenes.default_scene(), 1, 1000, False, False),
]


def run(scene, seed, substeps, collisions, t_acc, dtype):
    pos, vel = scenes.spawn(scene, seed=seed, dtype=dtype)
    p = oracle.make_params(radius=scene.radius, width=scene.width, height=scene.height, grid_size=scene.grid_size)
    out = oracle.run(p, pos, vel, substeps, collisions=collisions, t_acc=t_acc,
                     capture_substep=substeps - 1 if collisions else -1, pairs_cap=16 * scene.count)
    return pos, vel, out


def digest(*arrays):
    h = hashlib.sha256()
    for a in arrays:
        h.update(np.ascontiguousarray(a).tobytes())
    return h.hexdigest()


def main():
    digests = {}
    for name, scene, seed, substeps, collisions, t_acc in CASES:
        for tag, dtype in (("f64", np.float64), ("f32", np.float32)):
            pos, vel, out = run(scene, seed, substeps, collisions, t_acc, dtype)
            pairs = out["pairs"] if collisions else np.zeros((0, 2), dtype=np.uint64)
            np.savez_compressed(os.path.join(HERE, f"{name}_{tag}.npz"), pos=pos, vel=vel, pairs=pairs.astype(np.uint32),
                                meta=np.array(json.dumps({"count": scene.count, "width": scene.width, "height": scene.height,
                                                          "seed": seed, "substeps": substeps, "collisions": collisions, "t_acc": t_acc,
                                                          "pairs_total": out["stats"]["pairs_total"]})))
    for name, scene, seed, substeps, collisions, t_acc in DEFAULT_CASES:
        for tag, dtype in (("f64", np.float64), ("f32", np.float32)):
            pos, vel, out = run(scene, seed, substeps, collisions, t_acc, dtype)
            digests[f"{name}_{tag}"] = {"sha256_pos_vel": digest(pos, vel), "pairs_total": out["stats"]["pairs_total"],
                                        "seed": seed, "substeps": substeps, "collisions": collisions, "t_acc": t_acc}
    with open(os.path.join(HERE, "default_scene_digests.json"), "w") as f:
        json.dump(digests, f, indent=1, sort_keys=True)
    print("wrote", sorted(os.listdir(HERE)))


if __name__ == "__main__":
    main()
