#!/usr/bin/env python
"""What the host link can carry: the bound under bench.py's `e2e` number.

Copies pinned host buffers to the device, back, and both at once (two streams, the link is full duplex),
with the process pinned to each NUMA node in turn when the box has more than one.  Prints one JSON line.
Run on the GPU box:  python tools/pcie_probe.py [--mib 1024]
Under torchrun (N processes, one GPU each) every rank copies AT THE SAME TIME (a gloo barrier before each timed copy) and
rank 0 prints the per-rank rates, their sum, and where the pinned pages live (NUMA node of the buffers, CPU affinity):
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P tools/pcie_probe.py
"""
import argparse
import json
import os
import subprocess

import torch


def node_cpus():
    nodes = {}
    base = "/sys/devices/system/node"
    if not os.path.isdir(base):
        return nodes
    for d in sorted(os.listdir(base)):
        if d.startswith("node") and d[4:].isdigit():
            try:
                txt = open(os.path.join(base, d, "cpulist")).read().strip()
            except OSError:
                continue
            cpus = set()
            for part in txt.split(","):
                if not part:
                    continue
                a, _, b = part.partition("-")
                cpus.update(range(int(a), int(b or a) + 1))
            if cpus:
                nodes[int(d[4:])] = cpus
    return nodes


def measure(mib, reps=5, device=0, sync=None):
    n = mib << 20
    dev = torch.device("cuda", device)
    h_in = torch.empty(n, dtype=torch.uint8).pin_memory()
    h_out = torch.empty(n, dtype=torch.uint8).pin_memory()
    h_in.fill_(1)
    h_out.fill_(2)                       # touch the pages from this CPU set
    d_a = torch.empty(n, dtype=torch.uint8, device=dev)
    d_b = torch.ones(n, dtype=torch.uint8, device=dev)
    s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()

    def timed(fn):
        best = 1e9
        for _ in range(reps):
            torch.cuda.synchronize()
            if sync is not None:
                sync()                   # all ranks start their copy together
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            fn()
            s1.synchronize(); s2.synchronize()
            e1.record()
            torch.cuda.synchronize()
            best = min(best, e0.elapsed_time(e1))
        return best

    # each copy runs on its own stream, ordered after the start event only; timed() waits for both streams
    def h2d():
        s1.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(s1):
            d_a.copy_(h_in, non_blocking=True)

    def d2h():
        s2.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(s2):
            h_out.copy_(d_b, non_blocking=True)

    def both():
        h2d(); d2h()

    gib = n / 1e9
    return {"h2d_GBps": round(gib / (timed(h2d) * 1e-3), 1), "d2h_GBps": round(gib / (timed(d2h) * 1e-3), 1),
            "duplex_each_GBps": round(gib / (timed(both) * 1e-3), 1)}


def numa_of_buffer(t):
    """NUMA node holding the first page of a pinned tensor (move_pages with a NULL target list), or None."""
    import ctypes
    try:
        lib = ctypes.CDLL(None)
        page = ctypes.c_void_p(t.data_ptr() & ~4095)
        status = ctypes.c_int(-1)
        rc = lib.syscall(279, 0, 1, ctypes.byref(page), None, ctypes.byref(status), 0)   # __NR_move_pages on x86-64
        return int(status.value) if rc == 0 else None
    except Exception:  # noqa: BLE001
        return None


def main_ranks(a):
    import torch.distributed as dist

    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dist.init_process_group("gloo")
    probe = torch.empty(1 << 20, dtype=torch.uint8).pin_memory()
    probe.fill_(1)
    aff = sorted(os.sched_getaffinity(0))
    mine = {"rank": rank, "together": measure(a.mib, reps=3, device=local, sync=dist.barrier),
            "cpu_affinity": [aff[0], aff[-1]], "n_affinity": len(aff), "pinned_buffer_numa_node": numa_of_buffer(probe)}
    parts = [None] * world
    dist.all_gather_object(parts, mine)
    if rank == 0:
        out = {"mib": a.mib, "ranks": world, "host_cpus": os.cpu_count(), "numa_nodes": sorted(node_cpus()), "per_rank": parts,
               "aggregate_GBps": {k: round(sum(p["together"][k] for p in parts), 1) for k in ("h2d_GBps", "d2h_GBps", "duplex_each_GBps")}}
        print(json.dumps(out))
    dist.barrier()
    dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--mib", type=int, default=1024)
    a = ap.parse_args()
    if int(os.environ.get("WORLD_SIZE", "1")) > 1:
        return main_ranks(a)
    out = {"mib": a.mib, "cpus": os.cpu_count()}
    try:
        out["topo"] = subprocess.run(["nvidia-smi", "topo", "-m"], capture_output=True, text=True, timeout=20).stdout.splitlines()[:4]
    except Exception as exc:  # noqa: BLE001
        out["topo"] = str(exc)
    all_cpus = os.sched_getaffinity(0)
    out["default"] = measure(a.mib)
    nodes = node_cpus()
    if len(nodes) > 1:
        for node, cpus in nodes.items():
            use = cpus & all_cpus
            if not use:
                continue
            os.sched_setaffinity(0, use)
            out[f"node{node}"] = measure(a.mib)
        os.sched_setaffinity(0, all_cpus)
    print(json.dumps(out))


if __name__ == "__main__":
    main()
