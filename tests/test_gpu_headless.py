"""The headless C++ driver (host/headless_main.cpp: main.zig's init + loop + key handling over the C ABI through the
core_cuda.hpp mirror) against an oracle replay of the same key script."""
import json
import os
import subprocess

import numpy as np
import pytest

from _util import assert_bit_exact, run_oracle
from particle_simulator_b200 import build, scenes

pytestmark = pytest.mark.gpu


def replay(frames, keys, scene, pos, vel):
    """src/main.zig:95-130 with the oracle as updatePhysics."""
    script = {}
    for tok in keys.split(","):
        name, at = tok.split("@")
        script.setdefault(int(at), []).append(name)
    pause = nxt = t_acc = quit_ = False
    stepped = 0
    for frame in range(frames):
        if quit_:
            break
        for k in script.get(frame, []):
            if k == "escape": quit_ = not quit_
            elif k == "space": pause = not pause
            elif k == "right": nxt = True
            elif k == "g": t_acc = not t_acc
        if not pause or nxt:
            pos, vel, _ = run_oracle(scene, pos, vel, scene.sub_steps, collisions=True, t_acc=t_acc)
            nxt = False
            stepped += 1
    return pos, vel, stepped


@pytest.mark.parametrize("precision", ["f64", "f32"])
def test_headless_driver_replays_main_zig(tmp_path, precision):
    exe = build.build_headless()
    assert exe and os.path.exists(exe)
    keys = "g@6,space@10,right@12,right@13,space@16,g@20,escape@27"
    dump = tmp_path / "state.bin"
    out = subprocess.run([exe, "--frames", "40", "--seed", "3", "--keys", keys, "--dump", str(dump), "--precision", precision],
                         capture_output=True, text=True, timeout=120)
    assert out.returncode == 0, out.stderr
    info = json.loads(out.stdout.strip().splitlines()[-1])
    raw = np.fromfile(dump, dtype=np.uint8)
    n = int(raw[:8].view(np.uint64)[0])
    assert n == 13000
    pos = raw[8:8 + n * 16].view(np.float64).reshape(n, 2)
    vel = raw[8 + n * 16:8 + 2 * n * 16].view(np.float64).reshape(n, 2)
    scene = scenes.default_scene()
    pos0, vel0 = scenes.spawn(scene, seed=3)          # the C++ Spawner uses the same counter-based generator
    dt = np.float64 if precision == "f64" else np.float32
    # an f32 handle rounds the caller's f64 arrays once at upload and hands back widened f32 values: the oracle
    # runs in the same type from the same rounded start
    want_p, want_v, stepped = replay(40, keys, scene, pos0.astype(dt), vel0.astype(dt))
    assert info["frames"] == stepped and info["substeps"] == stepped * 8
    assert_bit_exact(pos.copy(), want_p.astype(np.float64), "positions")
    assert_bit_exact(vel.copy(), want_v.astype(np.float64), "velocities")
