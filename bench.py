#!/usr/bin/env python
"""bench.py — particle-updates/s of the particle step on B200, with its HBM roofline.

Contract (one JSON line on stdout from rank 0):
    python bench.py --gpus N --steps K --warmup W            # this library (libparticles_cuda via the C ABI)
    python bench.py --impl reference --gpus N --steps K ...  # the reference's CPU path (C++ port of core.zig)

Workload (BASELINE.json configs[2], the configuration the roofline metric is quoted on):
    64 Mi particles per GPU, integrate + wall bounce, fp32 SoA, reference scene constants.
    A "step" is one Solver.updatePhysics frame = sub_steps (8) passes of the hot kernel, one launch each.
    N > 1: every rank owns its own 64 Mi-particle index shard (no data-path collective) -> weak scaling.
"value" is timed with the state resident in HBM (fp32; roofline.f64 = the f64 build); "e2e" is the same frame driven through
the C ABI from pinned HOST buffers laid out like core.zig's []@Vector(2,f64), on an f64 handle, H2D and D2H inside the timed
region (e2e.f32_host_f32_device: the fp32 pair; e2e.resident_pos_f32: state resident, positions read back as f32).

Every run also times the collision-bearing configurations and reports them under config.collide_* (headline numbers) and in
the trailing "collide" object (everything): BASELINE.json configs[4], ONE 256 Mi-particle collision scene on all N GPUs (the
single-GPU handle at N = 1; x-slabs over NCCL send/recv at N > 1, strong scaling, with the per-stage breakdown and a 1 Mi-particle
run through the same NCCL path compared with the CPU oracle bit for bit - the run fails if it differs); at N = 1 also
configs[3] (16 Mi particles) and main.zig's default scene as a settled pile under gravity against the CPU port.
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
    ap.add_argument("--watchdog", type=int, default=780, help="seconds after which a stuck run dumps its stacks and exits (0 = off)")
    ap.add_argument("--no-f64", action="store_true", help="skip the f64 build's roofline / end-to-end figures")
    ap.add_argument("--no-collide", action="store_true", help="skip the collision scenes (BASELINE.json configs[3], configs[4])")
    ap.add_argument("--collide-particles", type=int, default=256 << 20, help="particles of the configs[4] collision scene (whole scene, all GPUs)")
    ap.add_argument("--collide-frames", type=int, default=3, help="timed updatePhysics frames of each collision scene")
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


T_START = time.time()


def log(*a):
    print(f"[{time.time() - T_START:7.1f}s]", *a, file=sys.stderr, flush=True)


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
                    f"(BASELINE.json configs[2]); step = one updatePhysics frame = 8 sub-steps; value / roofline: {precision} arithmetic "
                    f"on {precision} SoA state (roofline.f64: the f64 build); e2e: f64 host arrays, f64 arithmetic (core.zig as written), "
                    f"e2e.f32_host_f32_device: f32 both sides; collision scenes: see config.collide_* and the trailing `collide` object",
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
    log(f"[rank {rank}] integrate: device-resident timing")
    timed(torch, dist, lambda: sim.step(1), W)
    l0 = sim.stats()["kernel_launches"]
    clocks.start()
    ms = timed(torch, dist, lambda: sim.step(1), K)
    launches = sim.stats()["kernel_launches"] - l0
    value = n_total * S * K / (ms * 1e-3)
    launch_ms = ms / max(launches, 1)
    achieved = n * BYTES_PER_UPDATE[prec] / (launch_ms * 1e-3) / 1e9
    peak, peak_src = hbm_peak()

    # ---- the f64 build of the same kernel (core.zig as written, src/core.zig:5): its own roofline figures
    roof64 = None
    log(f"[rank {rank}] integrate: f64 build and end-to-end legs")
    if prec == "f32" and not args.no_f64:
        with Simulation(scene, precision="f64", collisions=False, device=local) as s64:
            s64.set_stream(stream)
            s64.upload(hp, hv)
            timed(torch, dist, lambda: s64.step(1), W)
            k64 = max(3, min(K, 20))
            ms64 = timed(torch, dist, lambda: s64.step(1), k64)
            lm64 = ms64 / (k64 * S)
            a64 = n * BYTES_PER_UPDATE["f64"] / (lm64 * 1e-3) / 1e9
            roof64 = {"achieved": a64, "frac": a64 / peak, "launch_ms": lm64, "value": n_total * S * k64 / (ms64 * 1e-3), "unit": UNIT,
                      "kernel": "integrate_bounce_kernel<double>", "algorithmic_bytes_per_launch": n * BYTES_PER_UPDATE["f64"]}

            # ---- end to end through the C ABI from host buffers, like for like with core.zig: f64 host arrays
            # ([]@Vector(2,f64)), f64 arithmetic on the device, f64 results — bit-identical to the reference's step
            e2e = None
            if not args.no_e2e:
                def frame64():
                    # Solver.updatePhysics on host slices: H2D of this frame's inputs (positions + velocities, pinned host
                    # memory), the 8 sub-steps, D2H of the result (positions: what main.zig reads to draw)
                    s64.frame_host(hp, hv, ho, None, 1)

                ke = max(3, min(K, 10))
                timed(torch, dist, frame64, 2)
                ms_e = timed(torch, dist, frame64, ke)
                e2e = {"value": n_total * S * ke / (ms_e * 1e-3), "unit": UNIT,
                       "h2d_bytes_per_step": int(hp.nbytes + hv.nbytes), "d2h_bytes_per_step": int(ho.nbytes),
                       "ms_per_step": ms_e / ke, "steps": ke, "host_dtype": "f64", "device_dtype": "f64",
                       "api": "ps_frame_host(f64 pos_in, vel_in -> pos_out) on an f64 handle, pinned host buffers, "
                              "2 Mi-particle chunks over 4 streams, copy engines both ways",
                       "f64_device": {"value": n_total * S * ke / (ms_e * 1e-3), "ms_per_step": ms_e / ke}}
    else:
        e2e = None
    # the fp32 pair: f32 host arrays, f32 arithmetic (north_star's layout; half the bytes over the host link)
    if not args.no_e2e:
        h32 = [torch.empty((n, 2), dtype=torch.float32, pin_memory=True) for _ in range(3)]
        a32 = [t.numpy() for t in h32]
        a32[0][...] = hp
        a32[1][...] = hv
        ke = max(3, min(K, 10))
        timed(torch, dist, lambda: sim.frame_host(a32[0], a32[1], a32[2], None, 1), 2)
        ms32 = timed(torch, dist, lambda: sim.frame_host(a32[0], a32[1], a32[2], None, 1), ke)
        f32_pair = {"value": n_total * S * ke / (ms32 * 1e-3), "ms_per_step": ms32 / ke, "host_dtype": "f32", "device_dtype": "f32",
                    "h2d_bytes_per_step": int(a32[0].nbytes + a32[1].nbytes), "d2h_bytes_per_step": int(a32[2].nbytes),
                    "api": "ps_frame_host_f32 on the f32 handle"}
        # what main.zig's draw loop actually needs each frame (src/main.zig:102-107): the state stays on the GPU, the
        # step runs there, and only positions come back, as f32 (the f64 -> f32 cast of :104-105 done on the device)
        pos32 = a32[2]
        def resident_frame():
            sim.step(1)
            sim.download(pos32, None)
        timed(torch, dist, resident_frame, 2)
        ms_r = timed(torch, dist, resident_frame, ke)
        resident = {"value": n_total * S * ke / (ms_r * 1e-3), "ms_per_step": ms_r / ke, "h2d_bytes_per_step": 0,
                    "d2h_bytes_per_step": int(pos32.nbytes),
                    "note": "state resident on the GPU, one updatePhysics frame, positions read back as f32 (what main.zig draws from)"}
        del h32, a32, pos32
        if e2e is None:      # --precision f64 or --no-f64: the pair that matches the handle's own type
            e2e = {"value": f32_pair["value"], "unit": UNIT, "h2d_bytes_per_step": f32_pair["h2d_bytes_per_step"],
                   "d2h_bytes_per_step": f32_pair["d2h_bytes_per_step"], "ms_per_step": f32_pair["ms_per_step"], "steps": ke,
                   "host_dtype": "f32", "device_dtype": prec, "api": f32_pair["api"]}
        e2e["f32_host_f32_device"] = f32_pair
        e2e["resident_pos_f32"] = resident
        # the host link is the ceiling (tools/pcie_probe.py; profiles/): ~55 GB/s each way for one process, and the
        # processes of one box share the host's memory system, so the aggregate does not grow with N
        e2e["bound"] = "host link (PCIe 5 x16 per GPU; all ranks share one host memory system)"
        gbs = (e2e["h2d_bytes_per_step"] + e2e["d2h_bytes_per_step"]) / (e2e["ms_per_step"] * 1e-3) / 1e9
        e2e["link_gbs_per_rank"] = gbs
        e2e["link_gbs_aggregate"] = gbs * world
    clocks.stop()

    extras = {}
    if not args.no_extras and rank == 0 and world == 1:
        extras = run_extras(torch, sim, scene, hp, hv, n, S, prec)
    sim.close()
    del host_pos, host_vel, host_out, hp, hv, ho

    # ---- the collision-bearing configurations (BASELINE.json configs[3] and configs[4]); every rank takes part
    collide = None
    if not args.no_collide:
        log(f"[rank {rank}] collision scenes")
        collide = collide_block(torch, dist, args, rank, world, local)
        log(f"[rank {rank}] collision scenes done")

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
        config = workload_config(args, prec)
        if collide:
            # the headline numbers of the collision scenes again, where the driver keeps the whole object
            for key in ("collide_256Mi", "collide_16Mi", "collide_16Mi_f64"):
                if key in collide and "error" not in collide[key]:
                    c = collide[key]
                    config[key] = {k: c.get(k) for k in ("updates_per_s", "ms_per_substep", "pairs", "rounds", "n_gpus", "scaling",
                                                          "roofline_frac", "slab_parity", "nccl_ranks") if k in c}
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": W,
            "ms_per_step": ms / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": prec, "data": "synthetic",
            "config": config,
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": ncu_traffic(kernel), "traffic_source": "committed ncu --set full capture (profiles/roofline_traffic.json), not this run",
                         "kernel": kernel, "peak_source": peak_src,
                         "algorithmic_bytes_per_launch": n * BYTES_PER_UPDATE[prec], "launch_ms": launch_ms, "f64": roof64},
            "cpu_baseline": cpu,
            "e2e": e2e,
            "gpu_launches": int(launches),
            "clocks": clocks.summary(),
        }
        if extras:
            line["extra"] = extras
        if collide:
            line["collide"] = collide      # last key: the driver's stdout tail keeps the end of the line
        print(json.dumps(line), flush=True)
    if dist is not None:
        dist.destroy_process_group()
    return 0


# ------------------------------------------------------------------------------------------------
# collision scenes inside the default run
# ------------------------------------------------------------------------------------------------
def substep_algorithmic_bytes(n, cells, pairs, precision):
    """Algorithmic bytes of one collision sub-step: the sum over its kernels."""
    rec = 16 if precision == "f32" else 32                      # one record {x, y, u, w}
    # bin (record in, cell + arrival out), place (12 in, 4 out), move (record + 12 in, record + 4 out), detect (record in, ~4 out:
    # head words and first-list items of the particles that touch something), scans (6 per cell), resolve (two records and two
    # head words in and out + list items per executed pair)
    return n * ((rec + 8) + 16 + (2 * rec + 16) + (rec + 4)) + 6 * cells + pairs * (4 * rec + 36)


def scene_cells(scene):
    return (int(scene.width) // int(scene.grid_size)) * (int(scene.height) // int(scene.grid_size))


def time_collision_scene(torch, dist, sim, scene, frames, warm_frames, world, precision, local):
    """Warm up, then time `frames` updatePhysics frames of a collision handle (single-GPU or slab) with CUDA events
    on the launching stream, barrier + synchronize on both sides, max over ranks.  Returns the result block."""
    S = scene.sub_steps
    stream = torch.cuda.current_stream().cuda_stream
    sim.set_stream(stream)
    for _ in range(max(warm_frames, 1)):
        sim.step(1)
    sim.sync()
    clocks = ClockSampler(local)
    clocks.start()
    l0 = sim.stats()["kernel_launches"]

    def frame():
        sim.step(1)

    ms = timed(torch, dist, frame, frames)
    sim.sync()                                                   # surfaces any error flag raised by the last sub-step
    clocks.stop()
    st = sim.stats()
    pairs, rounds = st["pairs_last"], st["rounds_last"]
    if dist is not None:
        t = torch.tensor([pairs], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        pairs = int(t.item())
        t = torch.tensor([rounds], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        rounds = int(t.item())
    n = scene.count
    ms_sub = ms / (frames * S)
    per_substep = substep_algorithmic_bytes(n, scene_cells(scene), pairs, precision)
    peak, peak_src = hbm_peak()
    achieved = per_substep / (ms_sub * 1e-3) / 1e9 / world      # per GPU
    return {"updates_per_s": n * S * frames / (ms * 1e-3), "ms_per_substep": ms_sub, "ms_per_frame": ms / frames,
            "pairs": pairs, "rounds": rounds, "n_gpus": world, "particles": n, "frames": frames, "warm_frames": warm_frames,
            "dtype": precision, "gpu_launches": int(st["kernel_launches"] - l0),
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": None,
                         "kernel": "whole collision sub-step (bin, scans, place, move, detect, resolve)", "peak_source": peak_src,
                         "algorithmic_bytes_per_substep": per_substep,
                         "note": "per GPU; detect and resolve are latency / sector bound, not streaming kernels (DESIGN.md section 4)"},
            "roofline_frac": achieved / peak, "clocks": clocks.summary(),
            "timing": "CUDA events on the launching stream around the frames, barrier + synchronize on both sides, max over ranks"}


def collide_block(torch, dist, args, rank, world, local):
    """BASELINE.json configs[4] (one 256 Mi-particle collision scene on all N GPUs: the single-GPU handle at N = 1,
    x-slabs with NCCL send/recv of migrants, halo and boundary states at N > 1; strong scaling) and, at N = 1,
    configs[3] (16 Mi particles).  At N > 1 a 1 Mi-particle scene also runs 24 sub-steps through the same NCCL path
    and is compared with the CPU oracle bit for bit: the run fails if it differs."""
    from particle_simulator_b200 import Simulation, SlabSimulation, scenes

    prec = "f32"
    out = {"precision": prec, "scaling": "strong"}
    frames = max(1, args.collide_frames)
    gloo = dist.new_group(backend="gloo") if world > 1 else None     # host-side object gathers (results, stage times)

    def new_slab(scene, contacts=0):
        idt = torch.zeros(128, dtype=torch.uint8, device="cuda")
        if rank == 0:
            idt = torch.frombuffer(bytearray(SlabSimulation.unique_id()), dtype=torch.uint8).clone().cuda()
        dist.broadcast(idt, 0)
        return SlabSimulation(scene, rank, world, bytes(idt.cpu().numpy().tobytes()), precision=prec, device=local,
                              contacts_per_particle=contacts)

    # ---- configs[4]
    n_big = args.collide_particles
    scene = scenes.scaled_scene(n_big)
    key = "collide_256Mi" if n_big == 256 << 20 else f"collide_{n_big >> 20}Mi"
    try:
        sim = new_slab(scene) if world > 1 else Simulation(scene, precision=prec, collisions=True, device=local)
        log(f"[rank {rank}] {key}: handle created")
        try:
            sim.spawn(1)
            log(f"[rank {rank}] {key}: spawned")
            blk = time_collision_scene(torch, dist, sim, scene, frames, 2, world, prec, local)
            blk["scaling"] = "strong"
            blk["workload"] = (f"{n_big >> 20}Mi particles, one scene on {world} GPU(s), full updatePhysics with uniform-grid collisions "
                               f"(BASELINE.json configs[4]), fp32; " +
                               ("x-slabs of whole cell columns, NCCL send/recv of migrants + halo + boundary states" if world > 1
                                else "single-GPU handle (ps_step)"))
            if world > 1:
                blk["nccl_ranks"] = world
                blk["path"] = "ps_slab_* (csrc/ps_slab.cuh) over ncclSend/ncclRecv"
                # per-stage breakdown: a second, shorter run with the stage events switched on (one more synchronisation per
                # sub-step, so it is not the run that is timed above)
                sim.set_timing(True)
                sim.step(3)
                sim.sync()
                tm = sim.timing()
                sim.set_timing(False)
                parts = [None] * world
                dist.all_gather_object(parts, tm, group=gloo)
                blk["stage_us_per_substep"] = {"rank0": parts[0], "max_over_ranks": {k: max(p[k] for p in parts) for k in parts[0]},
                                               "note": "PS_SLAB_TIMING stages, CUDA events, 16 sub-steps after 8 discarded"}
            out[key] = blk
        finally:
            sim.close()
    except Exception as e:  # noqa: BLE001
        out[key] = {"error": f"{type(e).__name__}: {e}"}
        log(f"[rank {rank}] collision scene failed: {e}")

    # ---- N > 1: the NCCL path against the oracle, bit for bit
    if world > 1:
        log(f"[rank {rank}] slab parity check")
        out["slab_parity"] = slab_parity(torch, dist, gloo, rank, world, local, new_slab, prec)
        if key in out and "error" not in out[key]:
            out[key]["slab_parity"] = out["slab_parity"]["result"]
        if out["slab_parity"]["result"] != "bit-exact":
            raise SystemExit(f"bench.py: slab path differs from the oracle: {out['slab_parity']}")

    # ---- configs[0] after pressing `g` (src/main.zig:124): the shipped 13 000-particle scene settled into a pile under
    # gravity, the deep-dependency-chain case of the exact-order resolve; frames 100-125, f64, against the CPU port
    if world == 1:
        try:
            out["gravity_pile_13k"] = gravity_pile(torch, local)
        except Exception as e:  # noqa: BLE001
            out["gravity_pile_13k"] = {"error": f"{type(e).__name__}: {e}"}

    # ---- configs[3] (one GPU)
    if world == 1:
        try:
            cs = scenes.scaled_scene(16 << 20)
            with Simulation(cs, precision=prec, collisions=True, device=local) as c:
                c.spawn(1)
                blk = time_collision_scene(torch, None, c, cs, frames, 2, 1, prec, local)
                blk["workload"] = "16Mi particles, full updatePhysics with uniform-grid collisions (BASELINE.json configs[3]), fp32, one GPU"
                out["collide_16Mi"] = blk
            # the same scene in core.zig's own arithmetic (f64 records: twice the bytes)
            with Simulation(cs, precision="f64", collisions=True, device=local) as c:
                c.spawn(1)
                blk = time_collision_scene(torch, None, c, cs, frames, 2, 1, "f64", local)
                blk["workload"] = "16Mi particles, full updatePhysics with collisions, f64 (core.zig as written), one GPU"
                out["collide_16Mi_f64"] = blk
        except Exception as e:  # noqa: BLE001
            out.setdefault("collide_16Mi", {"error": f"{type(e).__name__}: {e}"})
            out.setdefault("collide_16Mi_f64", {"error": f"{type(e).__name__}: {e}"})
        # like for like on the collision path: 1 Mi particles, 24 sub-steps, f64, the CPU port (one thread) against one GPU
        try:
            out["cpu_vs_gpu_collide_1Mi_f64"] = cpu_vs_gpu_collide(torch, local)
        except Exception as e:  # noqa: BLE001
            out["cpu_vs_gpu_collide_1Mi_f64"] = {"error": f"{type(e).__name__}: {e}"}
    return out


def cpu_vs_gpu_collide(torch, local, n=1 << 20, substeps=24):
    """BASELINE.md run B4 inside the line: full updatePhysics with collisions, f64, oracle/core_ref.hpp on one thread against
    libparticles_cuda (device-resident), same initial state, final states compared bit for bit."""
    import oracle   # checker + CPU baseline of this case only
    from particle_simulator_b200 import Simulation, scenes

    sc = scenes.scaled_scene(n)
    pos0, vel0 = scenes.spawn(sc, seed=1, dtype=np.float64)
    p, v = pos0.copy(), vel0.copy()
    params = oracle.make_params(radius=sc.radius, width=sc.width, height=sc.height, grid_size=sc.grid_size)
    secs = oracle.run(params, p, v, substeps, collisions=True)["seconds"]
    with Simulation(sc, precision="f64", collisions=True, device=local) as c:
        c.set_stream(torch.cuda.current_stream().cuda_stream)
        c.upload(pos0, vel0)
        c.step_substeps(8)                      # warm-up
        c.upload(pos0, vel0)
        ms = timed(torch, None, lambda: c.step_substeps(substeps), 1)
        gp, gv = c.download(dtype=np.float64)
    same = bool(np.array_equal(gp.view(np.uint8), p.view(np.uint8)) and np.array_equal(gv.view(np.uint8), v.view(np.uint8)))
    return {"particles": n, "substeps": substeps, "dtype": "f64", "cpu_port_updates_per_s": n * substeps / secs, "cpu_cores": 1,
            "gpu_updates_per_s": n * substeps / (ms * 1e-3), "speedup_vs_cpu_port": secs * 1e3 / ms, "bit_identical": same}


def gravity_pile(torch, local, settle_frames=100, timed_frames=25):
    """The default main.zig scene with gravity on (key `g`): settle, then time frames 100-125 on the GPU (f64, one
    handle, device-resident) and on the CPU port (single thread) from the same state; final states compared bit for bit."""
    import oracle   # checker + CPU baseline of this case only
    from particle_simulator_b200 import Simulation, scenes

    sc = scenes.default_scene()
    S = sc.sub_steps
    pos0, vel0 = scenes.spawn(sc, seed=1, dtype=np.float64)
    p, v = pos0.copy(), vel0.copy()
    params = oracle.make_params()
    oracle.run(params, p, v, settle_frames * S, collisions=True, t_acc=True)
    cpu = oracle.run(params, p, v, timed_frames * S, collisions=True, t_acc=True)
    with Simulation(sc, precision="f64", collisions=True, device=local) as c:
        c.set_stream(torch.cuda.current_stream().cuda_stream)
        c.upload(pos0, vel0)
        c.set_acc(True)
        c.step(settle_frames)
        c.sync()
        r0 = c.stats()["rounds_total"]
        ms = timed(torch, None, lambda: c.step(timed_frames), 1)
        st = c.stats()
        gp, gv = c.download(dtype=np.float64)
    n_up = sc.count * S * timed_frames
    same = bool(np.array_equal(gp.view(np.uint8), p.view(np.uint8)) and np.array_equal(gv.view(np.uint8), v.view(np.uint8)))
    return {"gpu_updates_per_s": n_up / (ms * 1e-3), "gpu_ms_per_substep": ms / (S * timed_frames),
            "cpu_port_updates_per_s": n_up / cpu["seconds"], "cpu_ms_per_substep": 1e3 * cpu["seconds"] / (S * timed_frames),
            "speedup_vs_cpu_port": (cpu["seconds"] * 1e3) / ms, "bit_identical": same,
            "rounds_per_substep": (st["rounds_total"] - r0) / (S * timed_frames), "pairs_last": st["pairs_last"], "dtype": "f64",
            "workload": f"default scene (13 000 particles, 720x640) with gravity on, frames {settle_frames}-{settle_frames + timed_frames} "
                        f"(settled pile), full updatePhysics; CPU = oracle/core_ref.hpp, one thread"}


def slab_parity(torch, dist, g, rank, world, local, new_slab, prec, n=1 << 20, substeps=24):
    """1 Mi particles, 24 sub-steps through ps_slab_* over NCCL on all ranks, gathered on rank 0 and compared with the
    CPU oracle (the checker, outside every timed region): positions, velocities and the pair total, bit for bit."""
    from particle_simulator_b200 import scenes

    scene = scenes.scaled_scene(n)
    sim = new_slab(scene, contacts=24)
    try:
        sim.spawn(1)
        sim.step_substeps(substeps)
        sim.sync()
        ids, p, v = sim.download_owned()
        st = sim.stats()
    finally:
        sim.close()
    parts = [None] * world
    dist.all_gather_object(parts, (ids, p, v, st["pairs_total"]), group=g)
    res = [None]
    if rank == 0:
        import oracle   # checker only

        dtype = np.float32 if prec == "f32" else np.float64
        pos0, vel0 = scenes.spawn(scene, seed=1)
        wp, wv = pos0.astype(dtype), vel0.astype(dtype)
        params = oracle.make_params(radius=scene.radius, width=scene.width, height=scene.height, grid_size=scene.grid_size)
        o = oracle.run(params, wp, wv, substeps, collisions=True)
        gp = np.full((n, 2), np.nan)
        gv = np.full((n, 2), np.nan)
        seen = np.zeros(n, dtype=np.int64)
        for i_, p_, v_, _ in parts:
            ii = i_.astype(np.int64)
            gp[ii] = p_
            gv[ii] = v_
            np.add.at(seen, ii, 1)
        pairs = int(sum(x[3] for x in parts))
        ok = bool((seen == 1).all() and np.array_equal(gp.astype(dtype).view(np.uint8), wp.view(np.uint8)) and
                  np.array_equal(gv.astype(dtype).view(np.uint8), wv.view(np.uint8)) and pairs == o["stats"]["pairs_total"])
        res[0] = {"result": "bit-exact" if ok else "MISMATCH", "particles": n, "substeps": substeps, "ranks": world, "dtype": prec,
                  "pairs_total": pairs, "oracle_pairs_total": o["stats"]["pairs_total"],
                  "checked": "every particle owned once; positions, velocities bit-identical to oracle/core_ref.hpp; pair totals equal"}
    dist.broadcast_object_list(res, 0, group=g)
    return res[0]


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
    # A rank that waits for ever (a peer that died, a message that never comes) would hold the whole box until the
    # driver's limit: after --watchdog seconds every rank dumps its Python stacks to stderr and exits non-zero.
    import faulthandler

    if args.watchdog > 0:
        faulthandler.dump_traceback_later(args.watchdog, exit=True)
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
