#!/usr/bin/env python
"""What the host link can carry: the bound under bench.py's `e2e` number.

Copies pinned host buffers to the device, back, and both at once (two streams, the link is full duplex),
with the process pinned to each NUMA node in turn when the box has more than one.  Prints one JSON line.
Run on the GPU box:  python tools/pcie_probe.py [--mib 1024]
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


def measure(mib, reps=5):
    n = mib << 20
    dev = torch.device("cuda:0")
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


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--mib", type=int, default=1024)
    a = ap.parse_args()
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
