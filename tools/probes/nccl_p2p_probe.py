#!/usr/bin/env python
"""Developer probe: how long does one grouped send+recv exchange between two neighbours take over NCCL, by message size?
(The slab path exchanges fixed-size messages twice per sub-step.)
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port P tools/probes/nccl_p2p_probe.py"""
import json
import os

import torch
import torch.distributed as dist

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
peer = rank ^ 1
out = {}
for kib in (4, 32, 256, 1024, 2048, 8192):
    n = kib << 10
    a = torch.ones(n, dtype=torch.uint8, device="cuda")
    b = torch.empty(n, dtype=torch.uint8, device="cuda")

    def exch():
        ops = [dist.P2POp(dist.isend, a, peer), dist.P2POp(dist.irecv, b, peer)]
        for w in dist.batch_isend_irecv(ops):
            w.wait()

    for _ in range(5):
        exch()
    torch.cuda.synchronize()
    dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(50):
        exch()
    e1.record()
    torch.cuda.synchronize()
    out[f"{kib} KiB"] = round(e0.elapsed_time(e1) / 50 * 1e3, 1)
if rank == 0:
    print(json.dumps({"us_per_bidirectional_exchange": out}))
dist.destroy_process_group()
