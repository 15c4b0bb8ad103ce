"""A rank-by-rank model of the multi-GPU slab protocol (test infrastructure).

It mirrors, step for step, what particle-simulator_b200/csrc/ps_slab.cuh does on each GPU — the same
messages (M0 migrants + halo, M1 boundary states to the right, then ownership hand-over instead of a message back), the
same ownership / filter / blocking rules — with the oracle's f64 arithmetic, so that the PROTOCOL can
be checked on a CPU: R modelled ranks must reproduce the single-process oracle bit for bit.

Messages go through a `transport` object so that the same model runs in one process (all ranks in a
loop) or as real processes over torch.distributed (gloo).
"""
from __future__ import annotations

import math

import numpy as np

import oracle


def slab_bounds(cols: int, nranks: int, rank: int):
    """Own global column range [x0, x1) of `rank` (same formula as ps_slab_create)."""
    return (rank * cols) // nranks, ((rank + 1) * cols) // nranks


class RankModel:
    def __init__(self, scene, rank, nranks, t_acc=False):
        self.scene, self.rank, self.nranks, self.t_acc = scene, rank, nranks, t_acc
        self.params = oracle.make_params(radius=scene.radius, width=scene.width, height=scene.height, grid_size=scene.grid_size,
                                         target_fps=scene.target_fps, sub_steps=scene.sub_steps, g=scene.g,
                                         boundary_damping=scene.boundary_damping, particle_damping=scene.particle_damping)
        self.rows, self.cols = oracle.grid_shape(self.params)
        self.x0, self.x1 = slab_bounds(self.cols, nranks, rank)
        self.gl = 1 if rank > 0 else 0
        self.gr = 1 if rank < nranks - 1 else 0
        self.col_off = self.x0 - self.gl
        self.ncl = (self.x1 - self.x0) + self.gl + self.gr
        self.dt = (1.0 / scene.target_fps) / scene.sub_steps
        self.pos = np.zeros((0, 2)); self.vel = np.zeros((0, 2)); self.pid = np.zeros(0, dtype=np.int64)
        self.pairs_total = 0

    # ---- helpers
    def gcol(self, pos):
        gs = self.scene.grid_size
        return np.minimum(np.floor(pos[:, 0] / gs).astype(np.int64), self.cols - 1)

    def grow(self, pos):
        gs = self.scene.grid_size
        return np.minimum(np.floor(pos[:, 1] / gs).astype(np.int64), self.rows - 1)

    def load_global(self, pos, vel):
        gx = self.gcol(pos)
        m = (gx >= self.x0) & (gx < self.x1)
        self.pos, self.vel, self.pid = pos[m].copy(), vel[m].copy(), np.nonzero(m)[0].astype(np.int64)

    # ---- phase 1: integrate owned particles, build M0 messages
    def phase_integrate(self):
        p, v = self.pos.copy(), self.vel.copy()
        if len(p):
            oracle.run(self.params, p, v, 1, collisions=False, t_acc=self.t_acc)
        gx = self.gcol(p)
        # a particle may leave the slab by more than one column (integrate + the previous resolve's position
        # corrections), but never past the neighbouring slab.  Owned particles include the left ghost column of
        # the previous sub-step (handed over by the left rank) and exclude the last own column (handed to the right).
        lo = slab_bounds(self.cols, self.nranks, self.rank - 1)[0] if self.gl else 0
        hi = slab_bounds(self.cols, self.nranks, self.rank + 1)[1] if self.gr else self.cols
        assert ((gx >= lo) & (gx < hi)).all(), "particle crossed a whole slab in one sub-step"
        to_l = (gx <= self.x0) if self.gl else np.zeros(len(p), bool)       # migrants (< x0) + halo (x0)
        to_r = (gx >= self.x1 - 1) if self.gr else np.zeros(len(p), bool)   # migrants (>= x1) + halo (x1-1)
        keep = (gx >= self.col_off) & (gx < self.col_off + self.ncl)        # inside the local grid (own + ghost columns)
        self._own = (p[keep], v[keep], self.pid[keep], gx[keep])
        pack = lambda m: (p[m].copy(), v[m].copy(), self.pid[m].copy())
        return pack(to_l), pack(to_r)

    # ---- phase 2: merge received particles, sort, detect (filtered), mark blocked
    def phase_build(self, from_l, from_r):
        p, v, pid, gx = self._own
        parts = [(p, v, pid)]
        for msg in (from_l, from_r):
            if msg is not None:          # keep what lands in the local grid: migrants into own columns + the neighbour's halo
                g = self.gcol(msg[0])
                m = (g >= self.col_off) & (g < self.col_off + self.ncl)
                parts.append((msg[0][m], msg[1][m], msg[2][m]))
        pos = np.concatenate([a[0] for a in parts]); vel = np.concatenate([a[1] for a in parts]); ids = np.concatenate([a[2] for a in parts])
        gx, gy = self.gcol(pos), self.grow(pos)
        lx = gx - self.col_off
        assert ((lx >= 0) & (lx < self.ncl)).all()
        cell = lx * self.rows + gy
        order = np.lexsort((ids, cell))
        self.rec_p, self.rec_v, self.ids = pos[order].copy(), vel[order].copy(), ids[order].copy()
        self.cell = cell[order]
        n = len(order)
        ncell = self.ncl * self.rows
        self.start = np.searchsorted(self.cell, np.arange(ncell + 3))
        # owned after the sub-step: all but the last own column and the right ghost column, which the right rank keeps
        self.own_b, self.own_e = 0, (self.start[(self.ncl - 2) * self.rows] if self.gr else n)
        frozen = self.rec_p.copy()
        r2sq = (2 * self.scene.radius) ** 2
        rows, ncl, start = self.rows, self.ncl, self.start
        col_own = lambda c: self.gl <= c < ncl - self.gr
        lists = []

        def seg(a, b, e, o, centre, out):
            for q in range(b, e):
                dx = frozen[q, 0] - frozen[a, 0]; dy = frozen[q, 1] - frozen[a, 1]
                if dx * dx + dy * dy < r2sq:
                    out.append((q, o, centre))

        for a in range(n):
            c = self.cell[a]; cx = c // rows; gy_ = c - cx * rows
            up, down, left, right = gy_ > 0, gy_ + 1 < rows, cx > 0, cx + 1 < ncl
            a_own = col_own(cx)
            l_own = left and col_own(cx - 1)
            r_own = right and col_own(cx + 1)
            own_b, own_e = start[c], start[c + 1]
            out = []
            if l_own:                                           # a as neighbour of centres in the left column
                cl = c - rows
                if up: seg(a, start[cl - 1], start[cl], 3, 0, out)
                seg(a, start[cl], start[cl + 1], 2, 0, out)
            if a_own:
                if up: seg(a, start[c - 1], own_b, 1, 0, out)
                seg(a, own_b, a, 0, 0, out)
                seg(a, own_b, a, 0, 1, out); seg(a, a + 1, own_e, 0, 1, out)
                if down: seg(a, own_e, start[c + 2], 1, 1, out)
                if right:
                    cr = c + rows
                    seg(a, start[cr], start[cr + 1], 2, 1, out)
                    if down: seg(a, start[cr + 1], start[cr + 2], 3, 1, out)
                if left and down:
                    cl = c - rows; seg(a, start[cl + 1], start[cl + 2], 4, 1, out)
                seg(a, a + 1, own_e, 0, 0, out)
            if r_own and up:                                    # a as neighbour of centres in the right column
                cr = c + rows; seg(a, start[cr - 1], start[cr], 4, 0, out)
            lists.append(out)
        self.lists = lists
        self.head = [0] * n
        self.adv = [-1] * n
        self.round = 0
        nblk = self.start[2 * rows] if self.gl else 0          # left ghost column + first own column wait for M1
        self.blocked = [a < nblk for a in range(n)]
        self.nblk = nblk
        self.W = [a for a in range(n) if lists[a] and not self.blocked[a]]
        self.pairs_total += sum(1 for a in range(n) for e in lists[a] if e[2])

    def _resolve(self, i, j):
        P, V, t_acc, dt, r = self.rec_p, self.rec_v, self.t_acc, self.dt, self.scene.radius
        pix, piy = P[i]; pjx, pjy = P[j]; vix, viy = V[i]; vjx, vjy = V[j]
        ddx = pjx - pix; ddy = pjy - piy
        dist = math.sqrt(ddx * ddx + ddy * ddy)
        if dist < 1e-10:
            return
        nx = ddx / dist; ny = ddy / dist
        van = (vjx - vix) * nx + (vjy - viy) * ny
        kk = (-(1.0 + self.scene.particle_damping)) * van / 2
        overlap = 2.0 * r - dist; corr = overlap * 0.5
        if overlap > 0:
            if not t_acc:
                P[i, 0] -= corr * nx; P[i, 1] -= corr * ny; P[j, 0] += corr * nx; P[j, 1] += corr * ny
            else:
                pax = (10000.0 * nx) * overlap; pay = (10000.0 * ny) * overlap
                V[i, 0] -= pax * dt; V[i, 1] -= pay * dt; V[j, 0] += pax * dt; V[j, 1] += pay * dt
        V[i, 0] -= nx * kk; V[i, 1] -= ny * kk; V[j, 0] += nx * kk; V[j, 1] += ny * kk

    # ---- phases 3/5: rounds until the work list is empty
    def phase_resolve(self):
        lists, head, adv = self.lists, self.head, self.adv
        W = self.W
        while W:
            r = self.round
            Wn = []
            for a in W:
                if adv[a] == r: continue
                ia = head[a]
                if ia >= len(lists[a]): continue
                (b, o, ca) = lists[a][ia]
                if adv[b] == r or self.blocked[b]: continue
                ib = head[b]
                if ib >= len(lists[b]): continue
                (pa, ob, cb) = lists[b][ib]
                if not (pa == a and ob == o and cb != ca): continue
                if b < a and b in self._Wset(W): continue
                i, j = (b, a) if ca else (a, b)
                self._resolve(i, j)
                head[a] = ia + 1; adv[a] = r; head[b] = ib + 1; adv[b] = r
                if ia + 1 < len(lists[a]): Wn.append(a)
                if ib + 1 < len(lists[b]): Wn.append(b)
            W = Wn
            self.round += 1
        self.W = []

    def _Wset(self, W):
        if getattr(self, "_wcache", None) is not W:
            self._wcache, self._wset = W, set(W)
        return self._wset

    def right_complete(self):
        cR = self.start[(self.ncl - 2) * self.rows] if self.gr else len(self.ids)
        return all(self.head[a] >= len(self.lists[a]) for a in range(cR, len(self.ids)))

    def m1_out(self):      # last own column + right ghost column, slot order
        cR = self.start[(self.ncl - 2) * self.rows]
        return self.rec_p[cR:].copy(), self.rec_v[cR:].copy(), self.ids[cR:].copy()

    def m1_in(self, msg):  # becomes the state of left ghost column + first own column
        p, v, ids = msg
        assert np.array_equal(ids, self.ids[:self.nblk]), "M1 order mismatch"
        self.rec_p[:self.nblk] = p; self.rec_v[:self.nblk] = v
        for a in range(self.nblk):
            self.blocked[a] = False
        self.W = [a for a in range(self.nblk) if self.head[a] < len(self.lists[a])]

    def finish(self):
        b, e = self.own_b, self.own_e
        self.pos, self.vel, self.pid = self.rec_p[b:e].copy(), self.rec_v[b:e].copy(), self.ids[b:e].copy()


def substep(model: RankModel, transport):
    """One sub-step of one rank.  transport.sendrecv(dst_send, payload, src_recv) exchanges with neighbours."""
    r, R = model.rank, model.nranks
    to_l, to_r = model.phase_integrate()
    from_l, from_r = transport.exchange(r, to_l if r > 0 else None, to_r if r < R - 1 else None)        # M0
    model.phase_build(from_l, from_r)
    model.phase_resolve()                                                                              # phase A
    if r < R - 1 and model.right_complete():
        transport.send(r, r + 1, model.m1_out()); sent = True
    else:
        sent = r == R - 1
    if r > 0:
        model.m1_in(transport.recv(r, r - 1))                                                          # M1
        model.phase_resolve()                                                                          # phase B
    if not sent:
        assert model.right_complete()
        transport.send(r, r + 1, model.m1_out())
    model.finish()          # no third message: the right rank keeps the boundary column it has just finished
