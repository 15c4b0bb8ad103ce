"""CPU model of detect_kernel's list-ordered hit mask (csrc/ps_collide.cuh: test_flat, list_runs, list_mask, emit_list).

The kernel never materialises the reference's global pair list (src/core.zig:164-209): every slot derives its own
sub-list from the 3x3 block around its cell — hits as bits in column order, eight shift-and-mask runs into list order,
entry = run base + bit position.  This file restates that arithmetic in Python (32-bit wrap-around included) and checks
it against the oracle's ordered pair list, so the run table is guarded by the CPU test tier as well.
"""
from __future__ import annotations

import numpy as np
import pytest

import oracle
from particle_simulator_b200 import scenes

M32 = 0xFFFFFFFF
CENTRE = 1 << 31


def _grid(scene, pos):
    """SpatialHashGrid.init + computeGridIndex + the stable counting sort (src/core.zig:61-67, :103-121, :145-159)."""
    cols, rows = int(scene.width) // int(scene.grid_size), int(scene.height) // int(scene.grid_size)
    i = np.minimum(np.floor(pos[:, 0] / scene.grid_size).astype(np.int64), cols - 1)
    j = np.minimum(np.floor(pos[:, 1] / scene.grid_size).astype(np.int64), rows - 1)
    cell = i * rows + j
    order = np.argsort(cell, kind="stable")                      # slot -> particle index
    start = np.zeros(rows * cols + 1, dtype=np.int64)
    np.add.at(start, cell + 1, 1)
    start = np.cumsum(start)
    return rows, cols, cell[order], order, start


def _stencil(start, c, rows, cols):
    gx, gy = divmod(c, rows)
    up, down = gy > 0, gy + 1 < rows

    def column(cc):
        s1, s2 = start[cc], start[cc + 1]
        return (start[cc - 1] if up else s1), s1, s2, (start[cc + 2] if down else s2)

    m = column(c)
    l = column(c - rows) if gx > 0 else (0, 0, 0, 0)
    r = column(c + rows) if gx + 1 < cols else (0, 0, 0, 0)
    return l, m, r


def model_sublist(a, pos_slot, start, c, rows, cols, r2sq):
    """Entries (partner slot, role) of slot `a`, produced the way detect_kernel's 32-bit path produces them."""
    (l0, l1, l2, l3), (m0, m1, m2, m3), (r0, r1, r2, r3) = _stencil(start, c, rows, cols)
    nL, nM, nR = l3 - l0, m3 - m0, r3 - r0
    total = nL + nM + nR
    assert total + (m2 - m1) <= 32, "scene too dense for the 32-bit path this model follows"
    # test_flat: bit t = t-th candidate in column order
    hits = 0
    for t in range(total):
        q = t + (l0 if t < nL else (m0 - nL if t < nL + nM else r0 - nL - nM))
        dx, dy = pos_slot[q, 0] - pos_slot[a, 0], pos_slot[q, 1] - pos_slot[a, 1]
        if dx * dx + dy * dy < r2sq:
            hits |= 1 << t
    hits &= ~(1 << (nL + (a - m0)))
    # list_runs
    t1 = l2 - l0
    t2 = t1 + (a - m0)
    t3 = t2 + (a - m1)
    t4 = t3 + (m3 - a - 1)
    t5 = t4 + (r3 - r1)
    t6 = t5 + (l3 - l2)
    t7 = t6 + (m2 - a - 1)
    t8 = t7 + (r1 - r0)
    assert t8 <= 30

    def run(src, length, dst):
        return ((hits >> src) & ((1 << length) - 1)) << dst

    am = nL + (a - m0)
    lm = (run(0, t1, 0) | run(nL, t2 - t1, t1) | run(nL + (m1 - m0), t3 - t2, t2) | run(am + 1, t4 - t3, t3) |
          run(nL + nM + (r1 - r0), t5 - t4, t4) | run(t1, t6 - t5, t5) | run(am + 1, t7 - t6, t6) | run(nL + nM, t8 - t7, t7))
    # emit_list: entry = base of the run + bit position, modulo 2^32
    bases = [(l0, t1), ((m0 - t1) & M32, t2), ((m1 - t2 + CENTRE) & M32, t3), ((a + 1 - t3 + CENTRE) & M32, t4),
             ((r1 - t4 + CENTRE) & M32, t5), ((l2 - t5 + CENTRE) & M32, t6), ((a + 1 - t6) & M32, t7), ((r0 - t7) & M32, t8)]
    out = []
    p = 0
    while lm >> p:
        if (lm >> p) & 1:
            base = next(b for b, end in bases if p < end)
            e = (base + p) & M32
            out.append((e & ((1 << 28) - 1), e >> 31))
        p += 1
    return out, bin(lm).count("1")


def reference_sublists(pairs, slot_of, n):
    """Per slot: the pairs of the ordered list that mention it, in list order, as (partner slot, role)."""
    lists = [[] for _ in range(n)]
    for i, j in pairs:                                           # {particle_i = neighbour, particle_j = centre}, Q4
        si, sj = int(slot_of[i]), int(slot_of[j])
        lists[si].append((sj, 0))
        lists[sj].append((si, 1))
    return lists


@pytest.mark.parametrize("count,width,height,seed", [(700, 100.0, 80.0, 1), (400, 40.0, 120.0, 2), (70, 64.0, 8.0, 3), (50, 8.0, 64.0, 4)])
def test_list_ordered_mask_reproduces_the_reference_order(count, width, height, seed):
    scene = scenes.Scene(count=count, width=width, height=height)
    pos0, vel0 = scenes.spawn(scene, seed=seed)
    params = oracle.make_params(radius=scene.radius, width=scene.width, height=scene.height, grid_size=scene.grid_size)
    # positions at detection time: integrate + wall bounce of sub-step 0 (the same arithmetic with collisions off)
    pos1, vel1 = pos0.copy(), vel0.copy()
    oracle.run(params, pos1, vel1, 1, collisions=False)
    p, v = pos0.copy(), vel0.copy()
    out = oracle.run(params, p, v, 1, collisions=True, capture_substep=0, pairs_cap=count * 32)
    pairs = out["pairs"]
    assert out["pairs_count"] == len(pairs) > 0

    rows, cols, cell_slot, order, start = _grid(scene, pos1)
    slot_of = np.empty(count, dtype=np.int64)
    slot_of[order] = np.arange(count)
    pos_slot = pos1[order]
    want = reference_sublists(pairs.astype(np.int64), slot_of, count)
    r2sq = (2.0 * scene.radius) * (2.0 * scene.radius)
    entries = 0
    for a in range(count):
        got, cnt = model_sublist(a, pos_slot, start, int(cell_slot[a]), rows, cols, r2sq)
        assert cnt == len(got)
        assert got == want[a], f"slot {a} (cell {cell_slot[a]})"
        entries += cnt
    assert entries == 2 * len(pairs)                             # every pair sits in both of its particles' sub-lists


def test_first_work_list_rule():
    """detect_kernel lists a particle for round 0 only if its first partner has a higher slot (runs 3, 4, 6, 7):
    those runs must be exactly the ones whose slots lie above `a`."""
    scene = scenes.Scene(count=500, width=80.0, height=80.0)
    pos0, vel0 = scenes.spawn(scene, seed=7)
    rows, cols, cell_slot, order, start = _grid(scene, np.clip(pos0, scene.radius, None))
    pos_slot = np.clip(pos0, scene.radius, None)[order]
    r2sq = (2.0 * scene.radius) ** 2
    seen = 0
    for a in range(scene.count):
        got, _ = model_sublist(a, pos_slot, start, int(cell_slot[a]), rows, cols, r2sq)
        if not got:
            continue
        (l0, l1, l2, l3), (m0, m1, m2, m3), (r0, r1, r2, r3) = _stencil(start, int(cell_slot[a]), rows, cols)
        t3 = (l2 - l0) + (a - m0) + (a - m1)
        t5 = t3 + (m3 - a - 1) + (r3 - r1)
        t6 = t5 + (l3 - l2)
        partner, role = got[0]
        higher_by_rule = partner > a
        # list bit p of the first entry, from the run that holds (partner, role); the kernel's test is p in [t3, t5) or p >= t6
        runs = [(l0, l2, 0), (m0, a, 0), (m1, a, 1), (a + 1, m3, 1), (r1, r3, 1), (l2, l3, 1), (a + 1, m2, 0), (r0, r1, 0)]
        pos_in_list, p = 0, None
        for k, (b, e, ro) in enumerate(runs):
            if p is None and b <= partner < e and ro == role:
                p = pos_in_list + (partner - b)
            pos_in_list += max(e - b, 0)
        assert p is not None
        assert ((t3 <= p < t5) or p >= t6) == higher_by_rule
        seen += 1
    assert seen > 50
