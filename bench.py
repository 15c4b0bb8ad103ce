#!/usr/bin/env python
"""bench.py — particle-updates/s of the particle step on B200, with its HBM roofline.

Contract (one JSON line on stdout from rank 0):
    python bench.py --gpus N --steps K --warmup W            # this library (libparticles_cuda via the C ABI)
    python bench.py --impl reference --gpus N --steps K ...  # the reference's CPU path (C++ port of core.zig)

Workload (BASELINE.json configs[2], the configuration the roofline metric is quoted on):
    64 Mi particles per GPU, integrate + wall bounce, fp32 SoA, reference scene constants.
    A "step" is one Solver.updatePhysics frame = sub_steps (8) passes of the hot kernel, one launch each.
    N > 1: every rank owns its own 64 Mi-particle index shard (no data-path collective) -> weak scaling.
"value" is timed with the state resident in HBM; "e2e" is the same frame driven through the C ABI
from pinned HOST buffers laid out like core.zig's []@Vector(2,f64), H2D and D2H inside the timed region.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

METRIC = "particle-updates/sec"
UNIT = "particle-updates/s"
N_PER_GPU = 64 << 20
BYTES_PER_UPDATE = {"f32": 32, "f64": 64}  # read + write px, py, vx, vy
HBM_FALLBACK_GBS = 6650.0                  # /opt/skills/guides/B200_PROFILING.md fallback


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--particles", type=int, default=N_PER_GPU, help="particles per GPU")
    ap.add_argument("--precision", default="f32", choices=["f32", "f64"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-extras", action="store_true")
    ap.add_argument("--variant", type=int, default=0, help="integrate kernel variant (0 = auto)")
    ap.add_argument("--sweep", action="store_true", help="time every integrate kernel variant and print a table to stderr")
    ap.add_argument("--ref-sample", type=int, default=4 << 20, help="particles per step of the CPU reference arm")
    ap.add_argument("--ref-threads", type=int, default=0, help="host threads of the CPU reference arm (0 = all this process may use)")
    ap.add_argument("--workload", default="integrate", choices=["integrate", "collide"],
                    help="integrate = BASELINE.json configs[2] (default, the headline); collide = configs[4]: ONE 256 Mi-particle "
                         "collision scene slab-decomposed over the ranks (strong scaling)")
    return ap.parse_args()


def host_cpu_model():
    try:
        with open("/proc/cpuinfo") as f:
            for line in f:
                if line.lower().startswith("model name"):
                    return line.split(":", 1)[1].strip()
    except OSError:
        pass
    return "unknown"


def env_int(name, default):
    try:
        return int(os.environ.get(name, default))
    except ValueError:
        return default


def log(*a):
    print(*a, file=sys.stderr, flush=True)


def spawn_chunked(scene, seed, start, count, dtype, out_pos, out_vel):
    from particle_simulator_b200 import scenes

    chunk = 8 << 20
    for s in range(0, count, chunk):
        m = min(chunk, count - s)
        p, v = scenes.spawn(scene, seed=seed, start=start + s, count=m, dtype=dtype)
        out_pos[s:s + m] = p
        out_vel[s:s + m] = v


# ------------------------------------------------------------------------------------------------
# CPU reference arm / cpu_baseline: the C++ restatement of core.zig (oracle/), single thread.
# ------------------------------------------------------------------------------------------------
def cpu_reference_rate(scene, n_sample, frames, warm_frames=0, dtype=np.float64, seed=1, threads=1):
    """updates/s of the oracle port on `n_sample` particles, `frames` updatePhysics frames (integrate + bounce).

    core.zig is single-threaded.  With threads > 1 the particle range is cut into that many contiguous slices and the
    same port runs on every slice at once (integration-only particles are independent, so the result is the same);
    the rate is then taken from the host clock around all slices."""
    import oracle  # the only place bench.py touches oracle/: as the timed CPU baseline, never as the product
    from concurrent.futures import ThreadPoolExecutor

    sub = scene.with_(count=n_sample)
    pos = np.empty((n_sample, 2), dtype=dtype)
    vel = np.empty((n_sample, 2), dtype=dtype)
    spawn_chunked(scene, seed, 0, n_sample, dtype, pos, vel)
    # integration never touches the grid, but the port allocates it at init as SpatialHashGrid.init does (9 B per cell,
    # over a gigabyte for this world, once per call): one whole-world cell keeps that out of every slice
    params = oracle.make_params(radius=sub.radius, width=sub.width, height=sub.height, grid_size=max(sub.width, sub.height),
                                target_fps=sub.target_fps, sub_steps=sub.sub_steps, g=sub.g,
                                boundary_damping=sub.boundary_damping, particle_damping=sub.particle_damping)
    acc = np.zeros_like(pos)
    threads = max(1, min(int(threads), n_sample or 1))
    cuts = [n_sample * t // threads for t in range(threads + 1)]
    pool = ThreadPoolExecutor(threads) if threads > 1 else None

    def frame_run(substeps):
        if pool is None:
            return oracle.run(params, pos, vel, substeps, collisions=False, acc=acc)["seconds"]
        t0 = time.perf_counter()          # ctypes drops the GIL inside the call: the slices really run side by side
        list(pool.map(lambda t: oracle.run(params, pos[cuts[t]:cuts[t + 1]], vel[cuts[t]:cuts[t + 1]], substeps,
                                           collisions=False, acc=acc[cuts[t]:cuts[t + 1]]), range(threads)))
        return time.perf_counter() - t0

    if warm_frames:
        frame_run(warm_frames * sub.sub_steps)
    secs = [frame_run(sub.sub_steps) for _ in range(frames)]
    if pool is not None:
        pool.shutdown()
    total = float(sum(secs))
    return n_sample * sub.sub_steps * frames / total, total, secs


ALL_THREADS_NOTE = ("not the reference (BASELINE.md B5): core.zig is single-threaded; the same port with the particle range cut into "
                    "one contiguous slice per host thread, all slices running side by side")


def run_reference(args):
    rank = env_int("RANK", 0)
    if rank != 0:
        return 0
    from particle_simulator_b200 import scenes

    n_total = args.particles * args.gpus
    scene = scenes.scaled_scene(n_total)
    n_sample = min(args.ref_sample, args.particles)
    rate, total, secs = cpu_reference_rate(scene, n_sample, args.steps, warm_frames=args.warmup, dtype=np.float64)
    ms = 1e3 * total / max(args.steps, 1)
    threads = args.ref_threads if args.ref_threads > 0 else (len(os.sched_getaffinity(0)) or 1)
    many, _, _ = cpu_reference_rate(scene, n_sample, max(1, min(args.steps, 10)), warm_frames=1, dtype=np.float64, threads=threads)
    sample = (f"{n_sample} of the {args.particles} particles per step, one updatePhysics frame (8 sub-steps, integrate + "
              f"wall bounce, f64 as core.zig is written), single thread (core.zig has no threads)")
    line = {
        "impl": "reference",
        "metric": METRIC, "value": rate, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
        "data": "synthetic",
        "config": workload_config(args, "f64"),
        "cpu_baseline": {"value": rate, "unit": UNIT, "cores": 1, "kind": "port", "sample": sample,
                         "host_cpus": os.cpu_count(), "host_cpu_model": host_cpu_model(),
                         "all_host_threads": {"value": many, "cores": threads, "note": ALL_THREADS_NOTE}},
        "e2e": {"value": rate, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "note": "Zig toolchain absent: the reference arm is oracle/core_ref.hpp, a C++ restatement of src/core.zig "
                "(g++ -O2 -ffp-contract=off); parity with the Zig binary is unpinned",
    }
    print(json.dumps(line), flush=True)
    return 0


def workload_config(args, precision):
    return {
        "workload": f"{args.particles >> 20}Mi particles per GPU, integrate + wall bounce "
                    f"(BASELINE.json configs[2]); step = one updatePhysics frame = 8 sub-steps",
        "particles_per_gpu": args.particles, "sub_steps": 8, "scene": "reference constants, world at 0.444 particles/cell",
        "sharding": "particle index range per rank, no collective",
        "l2": f"state >= {args.particles * 16 >> 20} MiB per GPU (fp32 SoA), larger than the 126 MB L2: no flush needed",
    }


# ------------------------------------------------------------------------------------------------
# clocks
# ------------------------------------------------------------------------------------------------
class ClockSampler:
    """Samples SM clock and throttle reasons of one GPU through NVML while the timed regions run."""

    def __init__(self, index: int):
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._stop = threading.Event()
        self._thr = None
        try:
            import pynvml

            pynvml.nvmlInit()
            self.nv = pynvml
            # NVML indexes physical devices; honour CUDA_VISIBLE_DEVICES when it lists integers
            vis = os.environ.get("CUDA_VISIBLE_DEVICES")
            phys = index
            if vis:
                try:
                    phys = int(vis.split(",")[index])
                except (ValueError, IndexError):
                    phys = index
            self.h = pynvml.nvmlDeviceGetHandleByIndex(phys)
            self.max_mhz = float(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
        except Exception as e:  # pragma: no cover
            self.nv = None
            self.err = str(e)

    def _loop(self):
        nv = self.nv
        names = {
            getattr(nv, "nvmlClocksThrottleReasonHwSlowdown", 0x8): "hw_slowdown",
            getattr(nv, "nvmlClocksThrottleReasonHwThermalSlowdown", 0x40): "hw_thermal_slowdown",
            getattr(nv, "nvmlClocksThrottleReasonSwThermalSlowdown", 0x20): "sw_thermal_slowdown",
            getattr(nv, "nvmlClocksThrottleReasonSwPowerCap", 0x4): "sw_power_cap",
            getattr(nv, "nvmlClocksThrottleReasonHwPowerBrakeSlowdown", 0x80): "hw_power_brake_slowdown",
        }
        while not self._stop.is_set():
            try:
                self.samples.append(float(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)))
                r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for bit, name in names.items():
                    if r & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            self._stop.wait(0.02)

    def start(self):
        if self.nv is not None:
            self._thr = threading.Thread(target=self._loop, daemon=True)
            self._thr.start()

    def stop(self):
        self._stop.set()
        if self._thr is not None:
            self._thr.join()

    def summary(self):
        if self.nv is None or not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": [], "note": "NVML unavailable"}
        return {"sm_mhz": float(np.median(self.samples)), "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(self.samples)}


# ------------------------------------------------------------------------------------------------
# our arm
# ------------------------------------------------------------------------------------------------
def hbm_peak():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        with open(path) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return HBM_FALLBACK_GBS, "fallback (B200_PROFILING.md, MEASURED_PEAKS.json absent)"


def ncu_traffic(kernel_key):
    """DRAM bytes per launch of the dominant kernel from the committed ncu --set full capture, or None."""
    try:
        with open(os.path.join(ROOT, "profiles", "roofline_traffic.json")) as f:
            return json.load(f).get(kernel_key)
    except Exception:
        return None


def timed(torch, dist, fn, iters):
    """fn() x iters bracketed by barrier + synchronize, CUDA events on the launching stream, max over ranks (ms)."""
    if dist is not None:
        dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    if dist is not None:
        dist.barrier()
    ms = e0.elapsed_time(e1)
    if dist is not None:
        t = torch.tensor([ms], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    return ms


def run_ours(args):
    import torch

    from particle_simulator_b200 import Simulation, capi, scenes

    rank, world, local = env_int("RANK", 0), env_int("WORLD_SIZE", 1), env_int("LOCAL_RANK", 0)
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; libparticles_cuda has no CPU path (use --impl reference for the CPU arm)")
    capi.load(build_if_missing=False)
    torch.cuda.set_device(local)
    dist = None
    if world > 1:
        import torch.distributed as dist_mod

        dist_mod.init_process_group("nccl", device_id=torch.device("cuda", local))
        dist = dist_mod
    n = args.particles
    n_total = n * world
    scene_all = scenes.scaled_scene(n_total)          # one world, index-sharded across ranks
    scene = scene_all.with_(count=n)
    prec = args.precision
    np_dt = np.float32 if prec == "f32" else np.float64
    K, W, S = args.steps, max(args.warmup, 3), scene.sub_steps

    # host inputs in core.zig's layout ([]@Vector(2,f64)), pinned; this rank's shard of the global index range
    t0 = time.time()
    host_pos = torch.empty((n, 2), dtype=torch.float64, pin_memory=True)
    host_vel = torch.empty((n, 2), dtype=torch.float64, pin_memory=True)
    host_out = torch.empty((n, 2), dtype=torch.float64, pin_memory=True)
    hp, hv, ho = host_pos.numpy(), host_vel.numpy(), host_out.numpy()
    spawn_chunked(scene_all, 1, rank * n, n, np.float64, hp, hv)
    log(f"[rank {rank}] spawned {n} particles in {time.time() - t0:.1f}s")

    # a real (non-default) stream: the library queues its kernels on it and torch's events time it
    torch.cuda.set_stream(torch.cuda.Stream())
    stream = torch.cuda.current_stream().cuda_stream
    assert stream != 0
    sim = Simulation(scene, precision=prec, collisions=False, device=local)
    sim.set_stream(stream)
    sim.debug_set_integrate_variant(args.variant)
    sim.debug_set_transfer_mode(env_int("PS_BENCH_TRANSFER_MODE", 0))
    sim.upload(hp, hv)

    clocks = ClockSampler(local)

    if args.sweep and rank == 0:
        for var in (1, 2, 3, 4, 5, 6):
            sim.debug_set_integrate_variant(var)
            timed(torch, None, lambda: sim.step(1), 3)
            ms = timed(torch, None, lambda: sim.step(1), 20)
            gbs = n * BYTES_PER_UPDATE[prec] * S * 20 / (ms * 1e-3) / 1e9
            log(f"variant {var}: {ms / (20 * S):.4f} ms/launch  {gbs:.0f} GB/s")
        sim.debug_set_integrate_variant(args.variant)

    # ---- device-resident: value + roofline
    timed(torch, dist, lambda: sim.step(1), W)
    l0 = sim.stats()["kernel_launches"]
    clocks.start()
    ms = timed(torch, dist, lambda: sim.step(1), K)
    launches = sim.stats()["kernel_launches"] - l0
    value = n_total * S * K / (ms * 1e-3)
    launch_ms = ms / max(launches, 1)
    achieved = n * BYTES_PER_UPDATE[prec] / (launch_ms * 1e-3) / 1e9
    peak, peak_src = hbm_peak()

    # ---- end to end through the C ABI from host buffers
    e2e = None
    if not args.no_e2e:
        def frame():
            # Solver.updatePhysics on host slices: H2D of this frame's inputs (positions + velocities, pinned host memory
            # in core.zig's f64 layout), the 8 sub-steps, D2H of the result (positions: what main.zig reads to draw)
            sim.frame_host(hp, hv, ho, None, 1)

        ke = max(3, min(K, 10))
        timed(torch, dist, frame, 2)
        ms_e = timed(torch, dist, frame, ke)
        e2e = {"value": n_total * S * ke / (ms_e * 1e-3), "unit": UNIT,
               "h2d_bytes_per_step": int(hp.nbytes + hv.nbytes), "d2h_bytes_per_step": int(ho.nbytes),
               "ms_per_step": ms_e / ke, "steps": ke,
               "api": "ps_frame_host(f64 pos_in, vel_in -> pos_out), pinned host buffers, 2 Mi-particle chunks over 4 streams, copy engines both ways"}
    clocks.stop()

    extras = {}
    if not args.no_extras and rank == 0 and world == 1:
        extras = run_extras(torch, sim, scene, hp, hv, n, S, prec)
    sim.close()

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        # timed in a fresh process: the same code as `--impl reference`, away from this process's CUDA context, pinned
        # buffers and helper threads
        import subprocess

        n_s = min(n, 16 << 20)
        cmd = [sys.executable, os.path.abspath(__file__), "--impl", "reference", "--particles", str(n), "--ref-sample", str(n_s),
               "--steps", "4", "--warmup", "1", "--ref-threads", str(args.ref_threads)]
        try:
            res = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
            cpu = json.loads(res.stdout.strip().splitlines()[-1])["cpu_baseline"]
            cpu["sample"] = (f"{n_s} of the {n} particles, 4 updatePhysics frames (32 sub-steps) of integrate + wall bounce in f64 "
                             f"(core.zig as written), oracle/core_ref.hpp single thread, own process")
        except Exception as e:  # noqa: BLE001
            cpu = {"error": f"CPU baseline failed: {e}"}

    if dist is not None:
        dist.barrier()
    if rank == 0:
        kernel = f"integrate_bounce_kernel<{'float' if prec == 'f32' else 'double'}>"
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": W,
            "ms_per_step": ms / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": prec, "data": "synthetic",
            "config": workload_config(args, prec),
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": ncu_traffic(kernel), "kernel": kernel, "peak_source": peak_src,
                         "algorithmic_bytes_per_launch": n * BYTES_PER_UPDATE[prec], "launch_ms": launch_ms},
            "cpu_baseline": cpu,
            "e2e": e2e,
            "gpu_launches": int(launches),
            "clocks": clocks.summary(),
        }
        if extras:
            line["extra"] = extras
        print(json.dumps(line), flush=True)
    if dist is not None:
        dist.destroy_process_group()
    return 0


def run_extras(torch, sim, scene, hp, hv, n, S, prec):
    """Reported beside the headline, never instead of it: the register-fused integrate pass, the f64 build,
    and the collision scene of BASELINE.json configs[3]."""
    from particle_simulator_b200 import Simulation, scenes

    out = {}
    stream = torch.cuda.current_stream().cuda_stream
    try:
        with Simulation(scene, precision=prec, collisions=False, fuse_substeps=S) as fs:
            fs.set_stream(stream)
            fs.upload(hp, hv)
            timed(torch, None, lambda: fs.step(1), 3)
            ms = timed(torch, None, lambda: fs.step(1), 20)
            out["fused_frame"] = {"value": n * S * 20 / (ms * 1e-3), "unit": UNIT,
                                  "note": f"{S} sub-steps per pass in registers (bit-identical); bytes/update = {BYTES_PER_UPDATE[prec]}/{S}"}
    except Exception as e:
        out["fused_frame"] = {"error": str(e)}
    other = "f64" if prec == "f32" else "f32"
    try:
        with Simulation(scene, precision=other, collisions=False) as s2:
            s2.set_stream(stream)
            s2.upload(hp, hv)
            timed(torch, None, lambda: s2.step(1), 3)
            ms = timed(torch, None, lambda: s2.step(1), 20)
            gbs = n * BYTES_PER_UPDATE[other] * S * 20 / (ms * 1e-3) / 1e9
            out[f"integrate_{other}"] = {"value": n * S * 20 / (ms * 1e-3), "unit": UNIT, "hbm_gbs": gbs}
    except Exception as e:
        out[f"integrate_{other}"] = {"error": str(e)}
    # the same end-to-end frame with f32 host buffers (ps_frame_host_f32): half the bytes over the host link
    try:
        h32 = [torch.empty((n, 2), dtype=torch.float32, pin_memory=True) for _ in range(3)]
        a32 = [t.numpy() for t in h32]
        a32[0][...] = hp
        a32[1][...] = hv
        timed(torch, None, lambda: sim.frame_host(a32[0], a32[1], a32[2], None, 1), 2)
        ms = timed(torch, None, lambda: sim.frame_host(a32[0], a32[1], a32[2], None, 1), 5)
        out["e2e_f32_host"] = {"value": n * S * 5 / (ms * 1e-3), "unit": UNIT, "ms_per_step": ms / 5,
                               "h2d_bytes_per_step": int(a32[0].nbytes + a32[1].nbytes), "d2h_bytes_per_step": int(a32[2].nbytes),
                               "note": "ps_frame_host_f32: a host that keeps f32 arrays; the headline e2e uses core.zig's f64 layout"}
        del h32, a32
    except Exception as e:
        out["e2e_f32_host"] = {"error": str(e)}
    try:
        nc = 16 << 20
        cs = scenes.scaled_scene(nc)
        pos = np.empty((nc, 2), dtype=np.float32)
        vel = np.empty((nc, 2), dtype=np.float32)
        spawn_chunked(cs, 1, 0, nc, np.float32, pos, vel)
        with Simulation(cs, precision="f32", collisions=True) as c:
            c.set_stream(stream)
            c.upload(pos, vel)
            timed(torch, None, lambda: c.step(1), 2)
            ms = timed(torch, None, lambda: c.step(1), 3)
            st = c.stats()
            out["collide_16Mi_f32"] = {"value": nc * S * 3 / (ms * 1e-3), "unit": UNIT, "ms_per_substep": ms / (3 * S),
                                       "pairs_last": st["pairs_last"], "rounds_last": st["rounds_last"],
                                       "note": "BASELINE.json configs[3]: full updatePhysics with exact-order resolve"}
    except Exception as e:
        out["collide_16Mi_f32"] = {"error": str(e)}
    # BASELINE.md section 3, runs B1-B4: the CPU port (f64 unless noted, one thread) beside the same scene on the GPU
    try:
        import oracle

        d13, m1 = scenes.default_scene(), scenes.scaled_scene(1 << 20)
        plan = (("B1", "collide_default_13k_f64", d13, 1000, True, "f64"), ("B2", "integrate_default_13k_f64", d13, 1000, False, "f64"),
                ("B3", "integrate_1Mi_f64", m1, 1000, False, "f64"), ("B3", "integrate_1Mi_f32", m1, 1000, False, "f32"),
                ("B4", "collide_1Mi_f64", m1, 24, True, "f64"))
        for tag, name, cs, substeps, coll, pr in plan:
            pos, vel = scenes.spawn(cs, seed=1, dtype=np.float64 if pr == "f64" else np.float32)
            params = oracle.make_params(radius=cs.radius, width=cs.width, height=cs.height, grid_size=cs.grid_size)
            p, v = pos.copy(), vel.copy()
            secs = oracle.run(params, p, v, substeps, collisions=coll)["seconds"]
            with Simulation(cs, precision=pr, collisions=coll) as c:
                c.set_stream(stream)
                c.upload(pos, vel)
                c.step_substeps(8)
                c.upload(pos, vel)
                ms = timed(torch, None, lambda: c.step_substeps(substeps), 1)
                gp, gv = c.download(dtype=pos.dtype)
            out[name] = {"plan": tag, "cpu_port_updates_per_s": cs.count * substeps / secs, "cpu_wall_s": secs,
                         "gpu_updates_per_s": cs.count * substeps / (ms * 1e-3), "gpu_wall_s": ms * 1e-3, "substeps": substeps,
                         "bit_identical": bool(np.array_equal(gp.view(np.uint8), p.view(np.uint8)) and
                                               np.array_equal(gv.view(np.uint8), v.view(np.uint8))),
                         "note": ("full updatePhysics with collisions" if coll else "integrate + wall bounce only") +
                                 f", {pr}: oracle/core_ref.hpp single thread vs libparticles_cuda (device-resident)"}
    except Exception as e:
        out["cpu_compare"] = {"error": str(e)}
    return out


def run_collide(args):
    """BASELINE.json configs[3]/[4]: one collision scene of --particles particles (default 256 Mi) on N GPUs.
    N = 1: the single-GPU handle; N > 1: x-slabs, one process per GPU, NCCL send/recv of migrants, halo and
    boundary states.  Strong scaling: the scene is the same for every N."""
    import torch

    from particle_simulator_b200 import Simulation, SlabSimulation, capi, scenes

    rank, world, local = env_int("RANK", 0), env_int("WORLD_SIZE", 1), env_int("LOCAL_RANK", 0)
    capi.load(build_if_missing=False)
    torch.cuda.set_device(local)
    n = args.particles if args.particles != N_PER_GPU else 256 << 20
    scene = scenes.scaled_scene(n)
    K, W, S = args.steps, max(args.warmup, 3), scene.sub_steps
    dist = None
    if world > 1:
        import torch.distributed as dist_mod

        dist_mod.init_process_group("gloo")     # host-side channel only: NCCL id, barriers, max over ranks
        dist = dist_mod
        idt = torch.zeros(128, dtype=torch.uint8)
        if rank == 0:
            idt = torch.frombuffer(bytearray(SlabSimulation.unique_id()), dtype=torch.uint8).clone()
        dist.broadcast(idt, 0)
        sim = SlabSimulation(scene, rank, world, bytes(idt.numpy().tobytes()), precision=args.precision, device=local)
    else:
        sim = Simulation(scene, precision=args.precision, collisions=True, device=local)
    sim.spawn(1)
    clocks = ClockSampler(local)
    for _ in range(W):
        sim.step(1)
    sim.sync()
    if dist is not None:
        dist.barrier()
    torch.cuda.synchronize()
    clocks.start()
    t0 = time.perf_counter()
    for _ in range(K):
        sim.step(1)
    sim.sync()
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    clocks.stop()
    if dist is not None:
        t = torch.tensor([dt], dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dt = float(t.item())
    st = sim.stats()
    launches = st["kernel_launches"]
    sim.close()
    if rank == 0:
        # whole sub-step against the HBM roofline: algorithmic bytes of its kernels / time per sub-step
        rec = 16 if args.precision == "f32" else 32                      # one record {x, y, u, w}
        cells = (int(scene.width) // int(scene.grid_size)) * (int(scene.height) // int(scene.grid_size))
        pairs = st["pairs_last"] * world                                 # rank 0's last sub-step, all ranks alike
        per_substep = n * ((rec + 8) + 16 + (2 * rec + 16) + (rec + 12)) + 6 * cells + pairs * (4 * rec + 36)
        peak, peak_src = hbm_peak()
        achieved = per_substep / (dt / (K * S)) / 1e9 / world            # per GPU
        roofline = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": None,
                    "kernel": "whole collision sub-step (bin, scans, place, move, detect, resolve)", "peak_source": peak_src,
                    "algorithmic_bytes_per_substep": per_substep,
                    "note": "detect and resolve are latency / sector bound, not streaming kernels (DESIGN.md section 4); per GPU"}
        line = {"metric": METRIC, "value": n * S * K / dt, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": W,
                "ms_per_step": 1e3 * dt / K, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
                "dtype": args.precision, "data": "synthetic",
                "config": {"workload": f"{n >> 20}Mi particles, full updatePhysics with uniform-grid collisions (BASELINE.json configs[4]); "
                                       f"step = one frame = 8 sub-steps; x-slab decomposition, NCCL send/recv halo + migration",
                           "particles": n, "sub_steps": S, "timing": "host clock around K frames, sync on both sides, max over ranks"},
                "roofline": roofline,
                "gpu_launches": int(launches), "pairs_last_rank0": st["pairs_last"], "rounds_last_rank0": st["rounds_last"],
                "clocks": clocks.summary()}
        print(json.dumps(line), flush=True)
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()
    return 0


def main():
    args = parse_args()
    # The contract is ONE JSON line on stdout.  Libraries loaded later write there too (NCCL prints its version banner
    # to fd 1), so everything else is sent to stderr and only the JSON line goes to the real stdout.
    sys.stdout.flush()
    real_stdout = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    sys.stdout = real_stdout
    try:
        if args.impl == "reference":
            return run_reference(args)
        if args.workload == "collide":
            return run_collide(args)
        return run_ours(args)
    finally:
        real_stdout.flush()


if __name__ == "__main__":
    sys.exit(main())
