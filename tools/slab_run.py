#!/usr/bin/env python
"""Run one collision scene slab-decomposed across the GPUs of one box (one process per GPU), time it, and
optionally check it against the oracle.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
        tools/slab_run.py --particles 1048576 --substeps 24 [--precision f32] [--gravity] [--check] [--spawn-on-device]
"""
import argparse
import faulthandler
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import numpy as np  # noqa: E402
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

from particle_simulator_b200 import SlabSimulation, scenes  # noqa: E402


def log(rank, *a):
    print(f"[rank {rank} {time.strftime('%H:%M:%S')}]", *a, file=sys.stderr, flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--particles", type=int, default=1 << 20)
    ap.add_argument("--width", type=float, default=0.0)
    ap.add_argument("--height", type=float, default=0.0)
    ap.add_argument("--substeps", type=int, default=24)
    ap.add_argument("--warmup", type=int, default=8)
    ap.add_argument("--precision", default="f32")
    ap.add_argument("--gravity", action="store_true")
    ap.add_argument("--check", action="store_true", help="compare with the oracle (small scenes)")
    ap.add_argument("--spawn-on-device", action="store_true")
    ap.add_argument("--dense-band", action="store_true", help="non-uniform scene: the default 720x640 world with 80 %% of --particles in a band astride a slab boundary")
    ap.add_argument("--seed", type=int, default=1)
    ap.add_argument("--hang-dump", type=int, default=90)
    ap.add_argument("--verbose", action="store_true")
    args = ap.parse_args()
    rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
    faulthandler.dump_traceback_later(args.hang_dump, exit=True)
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("gloo")          # host-side channel: the NCCL id and the result gather only
    scene = scenes.scaled_scene(args.particles) if not args.width else scenes.Scene(count=args.particles, width=args.width, height=args.height)
    band = None
    if args.dense_band:
        scene = scenes.default_scene().with_(count=args.particles)
        rng = np.random.default_rng(args.seed)
        n = args.particles
        bp = np.empty((n, 2))
        bv = rng.uniform(-200.0, 200.0, size=(n, 2))
        in_band = rng.random(n) < 0.8
        bp[:, 0] = np.where(in_band, np.clip(rng.normal(180.0, 25.0, n), scene.radius, scene.width - scene.radius), rng.uniform(scene.radius, scene.width, n))
        bp[:, 1] = rng.uniform(scene.radius, scene.height, n)
        band = (np.ascontiguousarray(bp), np.ascontiguousarray(bv))
    idt = torch.zeros(128, dtype=torch.uint8)
    if world > 1:
        if rank == 0:
            idt = torch.frombuffer(bytearray(SlabSimulation.unique_id()), dtype=torch.uint8).clone()
        dist.broadcast(idt, 0)
    log(rank, "creating slab handle", scene)
    sim = SlabSimulation(scene, rank, world, bytes(idt.numpy().tobytes()) if world > 1 else None, precision=args.precision,
                         device=local, contacts_per_particle=96 if args.dense_band else 24)
    log(rank, "columns", sim.columns())
    pos0 = vel0 = None
    if band is not None:
        pos0, vel0 = band
        sim.upload_global(pos0, vel0)
    elif args.spawn_on_device:
        sim.spawn(args.seed)
    else:
        pos0, vel0 = scenes.spawn(scene, seed=args.seed)
        sim.upload_global(pos0, vel0)
    log(rank, "owned", sim.owned_count())
    sim.set_acc(args.gravity)
    if not args.check:
        for s in range(args.warmup):
            sim.step_substeps(1)
            if args.verbose:
                log(rank, "warm sub-step", s, "done")
        sim.sync()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for s in range(args.substeps):
        sim.step_substeps(1)
        if args.verbose:
            log(rank, "sub-step", s, "queued")
    sim.sync()
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    tt = torch.tensor([dt], dtype=torch.float64)
    if world > 1:
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
    st = sim.stats()
    log(rank, f"{args.substeps} sub-steps in {dt * 1e3:.2f} ms; owned {sim.owned_count()}; pairs_last {st['pairs_last']} rounds_last {st['rounds_last']}")
    ok = None
    if args.check:
        ids, p, v = sim.download_owned()
        if world > 1:
            parts = [None] * world
            dist.all_gather_object(parts, (ids, p, v, st["pairs_total"]))
        else:
            parts = [(ids, p, v, st["pairs_total"])]
        if rank == 0:
            import oracle

            dtype = np.float64 if args.precision == "f64" else np.float32
            if pos0 is None:
                pos0, vel0 = scenes.spawn(scene, seed=args.seed)
            wp, wv = pos0.astype(dtype), vel0.astype(dtype)
            params = oracle.make_params(radius=scene.radius, width=scene.width, height=scene.height, grid_size=scene.grid_size)
            out = oracle.run(params, wp, wv, args.substeps, collisions=True, t_acc=args.gravity)
            gp = np.full((scene.count, 2), np.nan); gv = np.full((scene.count, 2), np.nan)
            seen = np.zeros(scene.count, dtype=int)
            for i_, p_, v_, _ in parts:
                gp[i_.astype(np.int64)] = p_; gv[i_.astype(np.int64)] = v_; seen[i_.astype(np.int64)] += 1
            pairs = sum(x[3] for x in parts)
            okp = np.array_equal(gp.astype(dtype).view(np.uint8), wp.view(np.uint8))
            okv = np.array_equal(gv.astype(dtype).view(np.uint8), wv.view(np.uint8))
            ok = bool((seen == 1).all() and okp and okv and pairs == out["stats"]["pairs_total"])
            log(rank, f"CHECK owned-once={bool((seen == 1).all())} pos={okp} vel={okv} pairs {pairs} vs {out['stats']['pairs_total']}")
            if not okp:
                bad = np.nonzero((gp.astype(dtype) != wp).any(axis=1))[0]
                log(rank, "first differing particles", bad[:10], "of", bad.size)
    if rank == 0:
        print(json.dumps({"particles": scene.count, "n_gpus": world, "precision": args.precision, "substeps": args.substeps,
                          "ms_per_substep": float(tt.item()) * 1e3 / args.substeps,
                          "updates_per_s": scene.count * args.substeps / float(tt.item()), "check": ok,
                          "pairs_last_rank0": st["pairs_last"], "rounds_last_rank0": st["rounds_last"]}), flush=True)
    sim.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    faulthandler.cancel_dump_traceback_later()
    return 0 if ok in (None, True) else 1


if __name__ == "__main__":
    sys.exit(main())
