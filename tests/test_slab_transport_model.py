"""CPU model of the one-sided slab transport's buffer discipline (csrc/ps_slab.cuh: slab_push_kernel / slab_wait_kernel).

Every rank owns one receive block: M0-from-left, M0-from-right and M1 buffers, each DOUBLE-buffered by the parity of the
sub-step, and one flag word per buffer that the sender sets to the sub-step's sequence number after the payload.  A sender
never asks whether the receiver has consumed the previous content of a buffer: the claim is that the data dependencies of a
sub-step (a rank cannot finish sub-step s without its neighbours' messages of sub-step s) already guarantee that a buffer of
parity p is free again when sub-step s + 2 wants to write it.  Here the ranks are threads with random delays; a push asserts
that what it overwrites has been consumed, and every wait checks that it reads the sequence number it expects.  The same
schedule with SINGLE buffers must trip the assertion (the check has teeth)."""
import random
import threading
import time

import pytest


class Block:
    def __init__(self, nbuf):
        self.flag = {}       # (kind, parity) -> sequence number last written
        self.payload = {}    # (kind, parity) -> (seq, sender)
        self.consumed = {}   # (kind, parity) -> sequence number last consumed
        self.nbuf = nbuf
        self.lock = threading.Lock()


def run_ranks(nranks, substeps, nbuf, seed, late_prob=0.2):
    rng = random.Random(seed)
    delays = [[rng.random() * 0.002 for _ in range(6 * substeps)] for _ in range(nranks)]
    late = [[rng.random() < late_prob for _ in range(substeps)] for _ in range(nranks)]     # hand-over only after phase B
    blocks = [Block(nbuf) for _ in range(nranks)]
    errors = []

    def push(dst, kind, seq, sender):
        key = (kind, seq % nbuf)
        b = blocks[dst]
        with b.lock:
            old = b.payload.get(key)
            if old is not None and b.consumed.get(key, 0) != old[0]:
                errors.append(f"rank {sender} overwrites {kind} seq {old[0]} of rank {dst} with seq {seq} before it was consumed")
            b.payload[key] = (seq, sender)
            b.flag[key] = seq                         # (the kernel: payload, system-wide fence, then the flag)

    def wait(me, kind, seq, deadline):
        key = (kind, seq % nbuf)
        b = blocks[me]
        while True:
            with b.lock:
                if b.flag.get(key) == seq:
                    got = b.payload[key]
                    if got[0] != seq:
                        errors.append(f"rank {me} expected {kind} seq {seq}, buffer holds seq {got[0]}")
                    return
            if time.time() > deadline or errors:
                errors.append(f"rank {me} timed out waiting for {kind} seq {seq}")
                return
            time.sleep(0.0002)

    def consume(me, kind, seq):
        # the kernels that read the message have run (M0: bin_recv / place / move; M1: unpack)
        key = (kind, seq % nbuf)
        b = blocks[me]
        with b.lock:
            got = b.payload.get(key)
            if got is None or got[0] != seq:
                errors.append(f"rank {me} reads {kind} seq {seq} but the buffer holds {got}")
            b.consumed[key] = seq

    def rank(me):
        d = iter(delays[me])
        deadline = time.time() + 30
        for s in range(1, substeps + 1):
            time.sleep(next(d))                                        # bin
            if me > 0: push(me - 1, "m0_from_right", s, me)
            if me + 1 < nranks: push(me + 1, "m0_from_left", s, me)
            if me > 0: wait(me, "m0_from_left", s, deadline)
            if me + 1 < nranks: wait(me, "m0_from_right", s, deadline)
            time.sleep(next(d))                                        # sort (reads the M0 messages), detect
            if me > 0: consume(me, "m0_from_left", s)
            if me + 1 < nranks: consume(me, "m0_from_right", s)
            time.sleep(next(d))                                        # phase A
            is_late = late[me][s - 1]
            if me + 1 < nranks and not is_late: push(me + 1, "m1", s, me)
            if me > 0:
                wait(me, "m1", s, deadline)
                consume(me, "m1", s)
            time.sleep(next(d))                                        # phase B
            if me + 1 < nranks and is_late: push(me + 1, "m1", s, me)
            if errors:
                return

    threads = [threading.Thread(target=rank, args=(r,)) for r in range(nranks)]
    for t in threads: t.start()
    for t in threads: t.join()
    return errors


@pytest.mark.parametrize("nranks,seed", [(2, 1), (3, 2), (4, 3), (8, 4)])
def test_double_buffers_are_never_overwritten_early(nranks, seed):
    assert run_ranks(nranks, 40, nbuf=2, seed=seed) == []


def test_single_buffers_would_be():
    # with one buffer per kind a fast left rank overwrites its right neighbour's M0 before that neighbour has sorted it in
    hits = 0
    for seed in range(12):
        if any("overwrites" in e for e in run_ranks(3, 40, nbuf=1, seed=seed)):
            hits += 1
    assert hits > 0
