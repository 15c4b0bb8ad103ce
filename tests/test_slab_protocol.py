"""The multi-GPU slab protocol on CPU (no GPU needed).

`_slab_model.RankModel` mirrors the per-rank steps of csrc/ps_slab.cuh (migration + halo exchange M0,
ownership filter, blocked boundary particles, M1 rightward / M2 leftward state hand-over) with the
oracle's arithmetic.  R modelled ranks must reproduce the single-process oracle bit for bit.  The first
test runs the ranks as threads in one process; the second runs world_size = 2 as real processes over
torch.distributed (gloo), which is the N > 1 coverage the CPU tier can give."""
import os
import queue
import threading

import numpy as np
import pytest

import oracle
from _slab_model import RankModel, slab_bounds, substep
from particle_simulator_b200 import scenes


class ThreadTransport:
    def __init__(self, nranks):
        self.q = {(s, d): queue.Queue() for s in range(nranks) for d in range(nranks)}

    def send(self, src, dst, payload):
        self.q[(src, dst)].put(payload)

    def recv(self, me, src):
        return self.q[(src, me)].get(timeout=20)

    def exchange(self, me, to_l, to_r):
        if to_l is not None: self.send(me, me - 1, to_l)
        if to_r is not None: self.send(me, me + 1, to_r)
        return (self.recv(me, me - 1) if to_l is not None else None), (self.recv(me, me + 1) if to_r is not None else None)


def run_threads(scene, pos0, vel0, nranks, substeps, t_acc):
    tr = ThreadTransport(nranks)
    models = [RankModel(scene, r, nranks, t_acc) for r in range(nranks)]
    errors = []

    def work(m):
        try:
            m.load_global(pos0, vel0)
            for _ in range(substeps):
                substep(m, tr)
        except Exception as e:  # pragma: no cover
            errors.append(e)

    th = [threading.Thread(target=work, args=(m,)) for m in models]
    [t.start() for t in th]
    [t.join() for t in th]
    if errors:
        raise errors[0]
    pos, vel = np.empty_like(pos0), np.empty_like(vel0)
    seen = np.zeros(len(pos0), dtype=int)
    for m in models:
        pos[m.pid] = m.pos; vel[m.pid] = m.vel; seen[m.pid] += 1
    assert (seen == 1).all(), "every particle must be owned by exactly one rank"
    return pos, vel, sum(m.pairs_total for m in models)


def test_slab_bounds_partition_all_columns():
    for cols in (7, 180, 24576):
        for n in (1, 2, 3, 8):
            b = [slab_bounds(cols, n, r) for r in range(n)]
            assert b[0][0] == 0 and b[-1][1] == cols
            assert all(b[i][1] == b[i + 1][0] for i in range(n - 1))


@pytest.mark.parametrize("nranks,t_acc,seed", [(2, False, 1), (3, True, 2), (4, False, 3), (8, True, 4)])
def test_modelled_ranks_reproduce_the_oracle(nranks, t_acc, seed):
    # a small dense world, 32 columns: with 8 ranks every slab is four columns wide
    scene = scenes.Scene(count=450, width=128.0, height=36.0)
    pos0, vel0 = scenes.spawn(scene, seed=seed)
    substeps = 10
    want_p, want_v = pos0.copy(), vel0.copy()
    out = oracle.run(oracle.make_params(width=scene.width, height=scene.height), want_p, want_v, substeps, collisions=True, t_acc=t_acc)
    got_p, got_v, pairs = run_threads(scene, pos0, vel0, nranks, substeps, t_acc)
    assert pairs == out["stats"]["pairs_total"]
    assert np.array_equal(got_p.view(np.uint8), want_p.view(np.uint8))
    assert np.array_equal(got_v.view(np.uint8), want_v.view(np.uint8))


def _gloo_worker(rank, world, port, tmpdir):
    import torch
    import torch.distributed as dist

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)

    class GlooTransport:
        def send(self, src, dst, payload):
            p, v, ids = payload
            dist.send(torch.tensor([len(ids)], dtype=torch.int64), dst)
            if len(ids):
                dist.send(torch.from_numpy(np.concatenate([p, v], axis=1).copy()), dst)
                dist.send(torch.from_numpy(ids.astype(np.int64).copy()), dst)

        def recv(self, me, src):
            n = torch.zeros(1, dtype=torch.int64)
            dist.recv(n, src)
            n = int(n.item())
            pv = torch.zeros((n, 4), dtype=torch.float64); ids = torch.zeros(n, dtype=torch.int64)
            if n:
                dist.recv(pv, src); dist.recv(ids, src)
            pv = pv.numpy()
            return pv[:, :2].copy(), pv[:, 2:].copy(), ids.numpy().copy()

        def exchange(self, me, to_l, to_r):
            # ranks are ordered on a line: even ranks send first, odd ranks receive first (no deadlock with blocking p2p)
            out = [None, None]
            for phase in (0, 1):
                if (me % 2) == phase:
                    if to_l is not None: self.send(me, me - 1, to_l)
                    if to_r is not None: self.send(me, me + 1, to_r)
                else:
                    if to_l is not None: out[0] = self.recv(me, me - 1)
                    if to_r is not None: out[1] = self.recv(me, me + 1)
            return out[0], out[1]

    scene = scenes.Scene(count=400, width=64.0, height=64.0)
    pos0, vel0 = scenes.spawn(scene, seed=9)
    m = RankModel(scene, rank, world, t_acc=False)
    m.load_global(pos0, vel0)
    tr = GlooTransport()
    for _ in range(8):
        substep(m, tr)
    np.savez(os.path.join(tmpdir, f"rank{rank}.npz"), pos=m.pos, vel=m.vel, pid=m.pid)
    dist.barrier()
    dist.destroy_process_group()


def test_world_size_2_over_gloo(tmp_path):
    import torch.multiprocessing as mp

    port = 29500 + (os.getpid() % 2000)
    mp.spawn(_gloo_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    scene = scenes.Scene(count=400, width=64.0, height=64.0)
    pos0, vel0 = scenes.spawn(scene, seed=9)
    want_p, want_v = pos0.copy(), vel0.copy()
    oracle.run(oracle.make_params(width=64.0, height=64.0), want_p, want_v, 8, collisions=True)
    got_p, got_v = np.empty_like(pos0), np.empty_like(vel0)
    for r in range(2):
        z = np.load(tmp_path / f"rank{r}.npz")
        got_p[z["pid"]] = z["pos"]; got_v[z["pid"]] = z["vel"]
    assert np.array_equal(got_p.view(np.uint8), want_p.view(np.uint8))
    assert np.array_equal(got_v.view(np.uint8), want_v.view(np.uint8))
