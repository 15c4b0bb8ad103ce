"""CPU model of resolve_kernel's rounds (csrc/ps_collide.cuh: resolve_item, head words, work lists).

The reference resolves its ordered pair list front to back (src/core.zig:211-213).  The kernel instead runs, round
after round, every pair that heads BOTH of its particles' sub-lists, from work lists that hold only the particles whose
head moved in the previous round.  This file restates that protocol in Python — same head word (stamp, position, head entry),
same fixed chunk + overflow pool for the other entries, same tie-break, same rule for the first work list — executes the items of every round in a random order, and
checks that the state ends bit-identical to the sequential loop.  Arithmetic: oracle/core_ref.py's resolve_collision.
"""
from __future__ import annotations

import random

import numpy as np
import pytest

import oracle
from oracle import core_ref
from particle_simulator_b200 import scenes


def sublists(pairs, slot_of, n):
    """Per slot, in list order: (partner slot, role) with role 1 = this particle is the centre j of {i, j}."""
    lists = [[] for _ in range(n)]
    for i, j in pairs:
        si, sj = int(slot_of[i]), int(slot_of[j])
        lists[si].append((sj, 0))
        lists[sj].append((si, 1))
    return lists


LIST_CHUNK = 8          # csrc/ps_collide.cuh: entries 1..8 of a sub-list live in the slot's own chunk, entry 0 in its head word


def lay_out(lists):
    """detect_kernel's layout: per slot a chunk of LIST_CHUNK words, plus a pool for the tails of long sub-lists."""
    chunks, pool = [[None] * LIST_CHUNK for _ in lists], []
    for a, l in enumerate(lists):
        ovf = len(l) > LIST_CHUNK + 1
        if ovf:
            chunks[a][LIST_CHUNK - 1] = len(pool)                # index of the overflow run
        for j in range(1, len(l)):
            e = (l[j], j + 1 == len(l))                          # (entry, LAST)
            if not ovf or j - 1 < LIST_CHUNK - 1:
                chunks[a][j - 1] = e
            else:
                pool.append(e)
    return chunks, pool


def load_entry(chunks, pool, a, j, ovf):
    idx = j - 1
    if not ovf or idx < LIST_CHUNK - 1:
        return chunks[a][idx]
    return pool[chunks[a][LIST_CHUNK - 1] + idx - (LIST_CHUNK - 1)]


def run_rounds(lists, execute, rng):
    """The rounds of resolve_kernel.  execute(i_slot, j_slot) resolves one pair {i = neighbour, j = centre}.

    Per particle one head word [stamp, pos, live, ovf, last, entry]: a readiness test looks at the two head words only;
    the thread that runs the pair fetches each particle's next entry from that particle's chunk (or the pool)."""
    n = len(lists)
    chunks, pool = lay_out(lists)
    head = [[0, 0, bool(l), len(l) > LIST_CHUNK + 1, len(l) == 1, l[0] if l else None] for l in lists]
    # detect_kernel: on the first list only if the first partner has a higher slot (round 0 runs pairs from the lower slot)
    work = [(a, lists[a][0]) for a in range(n) if lists[a] and lists[a][0][0] > a]
    r, executed = 0, 0
    done = [0] * n
    while work:
        rng.shuffle(work)                                        # threads of one round run in no particular order
        nxt = []
        for a, ea in work:
            b, a_centre = ea
            if r == 0 and b < a:
                continue
            sa, pos_a, live_a, ovf_a, last_a, ent_a = head[a]
            sb, pos_b, live_b, ovf_b, last_b, ent_b = head[b]
            if sa == r + 1 or sb == r + 1 or not live_b:
                continue
            if ent_b != (a, 1 - a_centre):                       # not the same pair seen from the other side
                continue
            if b < a and sb == r:                                # both on this round's list: the lower slot runs it
                continue
            assert live_a and ent_a == ea and lists[a][pos_a] == ea and lists[b][pos_b] == ent_b
            na = None if last_a else load_entry(chunks, pool, a, pos_a + 1, ovf_a)
            nb = None if last_b else load_entry(chunks, pool, b, pos_b + 1, ovf_b)
            execute(b, a) if a_centre else execute(a, b)
            executed += 1
            # a finished sub-list's head word is NOT written back (nobody looks at that particle again in this sub-step,
            # and the tie-break rests on stamps of earlier rounds): the stale word must do no harm
            done[a] += 1
            done[b] += 1
            if na:
                head[a] = [r + 1, pos_a + 1, True, ovf_a, na[1], na[0]]
            if nb:
                head[b] = [r + 1, pos_b + 1, True, ovf_b, nb[1], nb[0]]
            if na:
                nxt.append((a, na[0]))
            if nb:
                nxt.append((b, nb[0]))
        work = nxt
        r += 1
    assert all(d == len(l) for d, l in zip(done, lists)), "some sub-list was not drained"
    return r, executed


def detection_state(scene, seed):
    """Positions / velocities at detection time of sub-step 0, the ordered pair list, and the oracle's result."""
    pos0, vel0 = scenes.spawn(scene, seed=seed)
    params = oracle.make_params(radius=scene.radius, width=scene.width, height=scene.height, grid_size=scene.grid_size)
    pos1, vel1 = pos0.copy(), vel0.copy()
    oracle.run(params, pos1, vel1, 1, collisions=False)
    p, v = pos0.copy(), vel0.copy()
    out = oracle.run(params, p, v, 1, collisions=True, capture_substep=0, pairs_cap=scene.count * 64)
    assert out["pairs_count"] == len(out["pairs"])
    return pos1, vel1, out["pairs"].astype(np.int64), p, v


def slot_order(scene, pos):
    cols, rows = int(scene.width) // int(scene.grid_size), int(scene.height) // int(scene.grid_size)
    i = np.minimum(np.floor(pos[:, 0] / scene.grid_size).astype(np.int64), cols - 1)
    j = np.minimum(np.floor(pos[:, 1] / scene.grid_size).astype(np.int64), rows - 1)
    order = np.argsort(i * rows + j, kind="stable")
    slot_of = np.empty(len(pos), dtype=np.int64)
    slot_of[order] = np.arange(len(pos))
    return order, slot_of


@pytest.mark.parametrize("count,width,height,t_acc,seed", [
    (600, 100.0, 80.0, False, 1),        # reference-like density: isolated pairs and short chains
    (350, 40.0, 40.0, False, 2),         # packed: long dependency chains, many same-cell pairs (listed twice, Q1)
    (350, 40.0, 40.0, True, 3),          # the penalty branch of resolve_collision (Q5)
    (120, 8.0, 200.0, False, 4),         # two cell columns
])
def test_rounds_equal_the_sequential_loop(count, width, height, t_acc, seed):
    scene = scenes.Scene(count=count, width=width, height=height)
    pos1, vel1, pairs, want_p, want_v = detection_state(scene, seed)
    assert len(pairs) > count // 8
    order, slot_of = slot_order(scene, pos1)
    solver = core_ref.Solver(scene.target_fps, scene.sub_steps, scene.g, scene.boundary_damping, scene.particle_damping)

    seq = core_ref.Particles(pos1, vel1, scene.radius)
    for i, j in pairs:                                            # src/core.zig:211-213
        core_ref.resolve_collision(seq, int(i), int(j), solver.dt, t_acc, solver.particle_damping)
    if not t_acc:                                                 # the C++ oracle ran this sub-step with gravity off
        assert np.array_equal(np.array(seq.positions), want_p) and np.array_equal(np.array(seq.velocities), want_v)

    lists = sublists(pairs, slot_of, count)
    for trial in range(3):
        par = core_ref.Particles(pos1, vel1, scene.radius)

        def execute(si, sj):
            core_ref.resolve_collision(par, int(order[si]), int(order[sj]), solver.dt, t_acc, solver.particle_damping)

        rounds, executed = run_rounds(lists, execute, random.Random(100 * seed + trial))
        assert executed == len(pairs)
        assert par.positions == seq.positions and par.velocities == seq.velocities, f"trial {trial}: state differs after {rounds} rounds"
        assert rounds >= 1
