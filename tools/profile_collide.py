#!/usr/bin/env python
"""Run the collision scene for a few frames (profiling target for ncu; also prints CUDA-event timings).

    python tools/profile_collide.py [--particles N] [--precision f32|f64] [--warm-frames W] [--frames F] [--gravity]
"""
import argparse
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import numpy as np  # noqa: E402
import torch  # noqa: E402

from particle_simulator_b200 import Simulation, scenes  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--particles", type=int, default=16 << 20)
ap.add_argument("--precision", default="f32")
ap.add_argument("--warm-frames", type=int, default=1)
ap.add_argument("--frames", type=int, default=1)
ap.add_argument("--gravity", action="store_true")
ap.add_argument("--spawn-on-device", action="store_true")
ap.add_argument("--default-scene", action="store_true", help="main.zig's 13 000-particle 720x640 scene instead of a scaled square world")
args = ap.parse_args()

scene = scenes.default_scene() if args.default_scene else scenes.scaled_scene(args.particles)
if args.default_scene:
    args.particles = scene.count
dt = np.float32 if args.precision == "f32" else np.float64
n = args.particles
if not args.spawn_on_device:
    pos = np.empty((n, 2), dtype=dt)
    vel = np.empty((n, 2), dtype=dt)
    chunk = 8 << 20
    for s in range(0, n, chunk):
        m = min(chunk, n - s)
        p, v = scenes.spawn(scene, seed=1, start=s, count=m, dtype=dt)
        pos[s:s + m] = p
        vel[s:s + m] = v
torch.cuda.set_stream(torch.cuda.Stream())
with Simulation(scene, precision=args.precision, collisions=True) as sim:
    sim.set_stream(torch.cuda.current_stream().cuda_stream)
    if args.spawn_on_device:
        sim.spawn(1)
    else:
        sim.upload(pos, vel)
    sim.set_acc(args.gravity)
    sim.step(args.warm_frames)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    sim.step(args.frames)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    st = sim.stats()
print(json.dumps({"particles": n, "precision": args.precision, "frames": args.frames, "ms_per_substep": ms / (args.frames * scene.sub_steps),
                  "updates_per_s": n * args.frames * scene.sub_steps / (ms * 1e-3), "pairs_last": st["pairs_last"],
                  "rounds_last": st["rounds_last"], "max_cell": st["max_cell_count"]}))
