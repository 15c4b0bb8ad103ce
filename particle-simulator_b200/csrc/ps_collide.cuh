// ps_collide.cuh — the collision half of updatePhysics (src/core.zig:262-266): uniform-grid
// broadphase, ordered contact detection and the exact-order resolve.
//
// The reference resolves an ordered pair list sequentially (Gauss-Seidel, src/core.zig:211-213);
// the result depends on that order, so this file reproduces the order instead of
// approximating it:
//
//   bin      integrate + bounce (ps_integrate.cuh arithmetic) fused with computeCoordinate /
//            computeGridIndex (:103-121) and the per-cell count (:124-134, one atomic per particle)
//   scan     computeStartArray (:137-142), ps_scan.cuh
//   scatter  computeSortedArray (:145-159).  Atomics hand out slots inside a cell in arbitrary
//            order, so a second kernel ranks each cell's few particles by index, which is the
//            order the reference's stable counting sort produces; the same kernel gathers
//            {px,py,vx,vy} into a cell-sorted array of 16/32-byte records ("slots").
//   detect   every slot walks its 3x3 cell block once in a FIXED segment order that visits its
//            contacts exactly in the order the reference's pair list (:164-209) mentions that
//            particle.  The position of an entry in the global list is the lexicographic key
//            (centre slot, offset index, neighbour slot), so each particle's own sub-list can
//            be produced without materialising or sorting the global list.
//   resolve  a pair may run as soon as it heads BOTH of its particles' sub-lists: pairs that
//            share no particle commute, so executing all such "ready" pairs together, round
//            after round, is bit-identical to the sequential loop.  One persistent cooperative
//            kernel runs the rounds (one grid barrier each) over work lists that hold only the
//            particles whose head changed in the previous round, so total work is O(pairs).
//   scatter-back  sorted records -> SoA arrays in original index order.
#pragma once
#include <cooperative_groups.h>

#include "ps_common.cuh"
#include "ps_integrate.cuh"
#include "ps_scan.cuh"

namespace ps {

namespace cg = cooperative_groups;

// ---- record types -------------------------------------------------------------------------
template <class Real> struct Rec;   // one particle: {px, py, vx, vy}
template <> struct Rec<float>  { float  x, y, u, w; };
template <> struct Rec<double> { double x, y, u, w; };

__device__ __forceinline__ Rec<float> ld_rec_cg(const Rec<float>* p) {
    const float4 v = __ldcg(reinterpret_cast<const float4*>(p));
    return Rec<float>{v.x, v.y, v.z, v.w};
}
__device__ __forceinline__ Rec<double> ld_rec_cg(const Rec<double>* p) {
    const double2 a = __ldcg(reinterpret_cast<const double2*>(p));
    const double2 b = __ldcg(reinterpret_cast<const double2*>(p) + 1);
    return Rec<double>{a.x, a.y, b.x, b.y};
}
__device__ __forceinline__ void st_rec(Rec<float>* p, const Rec<float>& r) {
    *reinterpret_cast<float4*>(p) = make_float4(r.x, r.y, r.u, r.w);
}
__device__ __forceinline__ void st_rec(Rec<double>* p, const Rec<double>& r) {
    reinterpret_cast<double2*>(p)[0] = make_double2(r.x, r.y);
    reinterpret_cast<double2*>(p)[1] = make_double2(r.u, r.w);
}
__device__ __forceinline__ float2  ld_pos(const Rec<float>* p)  { return *reinterpret_cast<const float2*>(p); }
__device__ __forceinline__ double2 ld_pos(const Rec<double>* p) { return *reinterpret_cast<const double2*>(p); }

__device__ __forceinline__ float  sqrt_rn(float x)  { return __fsqrt_rn(x); }
__device__ __forceinline__ double sqrt_rn(double x) { return __dsqrt_rn(x); }
__device__ __forceinline__ float  div_rn(float a, float b)   { return __fdiv_rn(a, b); }
__device__ __forceinline__ double div_rn(double a, double b) { return __ddiv_rn(a, b); }
__device__ __forceinline__ float  floor_r(float x)  { return floorf(x); }
__device__ __forceinline__ double floor_r(double x) { return floor(x); }

// ---- device-side bookkeeping ----------------------------------------------------------------
struct DevState {
    uint32_t flags;            // bit0: cell overflow (>255), bit1: contact capacity exceeded
    uint32_t max_cell;         // running maximum cell population
    uint32_t inc_total;        // contact-list entries of this sub-step (= 2 * pairs)
    uint32_t ctr_total;        // pairs of this sub-step
    uint32_t wl_count[3];      // rotating work-list lengths
    uint32_t rounds_last;
    unsigned long long pairs_total, rounds_total;
};
constexpr uint32_t FLAG_CELL_OVERFLOW = 1u, FLAG_CAPACITY = 2u;

// contact-list entry: partner slot (28 bits) | offset index (3 bits) | role (1 = this particle is the centre)
constexpr uint32_t ENT_SLOT_MASK = (1u << 28) - 1;
__device__ __forceinline__ uint32_t make_entry(uint32_t partner, uint32_t o, uint32_t centre) { return partner | (o << 28) | (centre << 31); }

// computeCoordinate + computeGridIndex (src/core.zig:103-121)
template <class Real>
__device__ __forceinline__ uint32_t cell_of(Real x, Real y, const StepConst<Real>& k) {
    const Real fi = floor_r(div_rn(x, k.grid_size));
    const Real fj = floor_r(div_rn(y, k.grid_size));
    // positions are inside [r, dim-r] here; the clamp mirrors @min(i, cols-1) and guards NaN / negatives
    uint32_t i = fi > (Real)0 ? (fi < (Real)k.cols ? (uint32_t)fi : k.cols - 1) : 0u;
    uint32_t j = fj > (Real)0 ? (fj < (Real)k.rows ? (uint32_t)fj : k.rows - 1) : 0u;
    return i * k.rows + j;
}

// ---- bin: integrate + bounce + cell key + per-cell count ------------------------------------
template <class Real>
__global__ void __launch_bounds__(256)
integrate_bin_kernel(Real* __restrict__ px, Real* __restrict__ py, Real* __restrict__ vx, Real* __restrict__ vy,
                     uint32_t n, StepConst<Real> k, uint32_t* __restrict__ cell_count,
                     uint32_t* __restrict__ cell_id, uint32_t* __restrict__ arrival, DevState* __restrict__ ds) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    if (ds->flags & FLAG_CAPACITY) return;                      // a failed sub-step freezes the state (see ps_status PS_ERR_CAPACITY)
    Real x = px[i], y = py[i], u = vx[i], w = vy[i];
    step_component<Real>(x, u, k.ax_dt, k.dt, k.lo, k.hi_x, k.neg_damp);
    step_component<Real>(y, w, k.ay_dt, k.dt, k.lo, k.hi_y, k.neg_damp);
    px[i] = x; py[i] = y; vx[i] = u; vy[i] = w;
    const uint32_t c = cell_of<Real>(x, y, k);
    const uint32_t a = atomicAdd(cell_count + c, 1u);           // grid_array[c] += 1   src/core.zig:132
    cell_id[i] = c;
    arrival[i] = a;
    if (a >= 255u) atomicOr(&ds->flags, FLAG_CELL_OVERFLOW);    // the reference's u8 would wrap here
}

// ---- scatter: unordered placement, then rank by particle index inside each cell --------------
__global__ void __launch_bounds__(256)
place_kernel(const uint32_t* __restrict__ cell_id, const uint32_t* __restrict__ arrival,
             const uint32_t* __restrict__ cell_start, uint32_t n, uint32_t* __restrict__ tmp_slot_id) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    tmp_slot_id[cell_start[cell_id[i]] + arrival[i]] = i;
}

template <class Real>
__global__ void __launch_bounds__(256)
rank_gather_kernel(const uint32_t* __restrict__ tmp_slot_id, const uint32_t* __restrict__ cell_id,
                   const uint32_t* __restrict__ cell_start, uint32_t n,
                   const Real* __restrict__ px, const Real* __restrict__ py, const Real* __restrict__ vx, const Real* __restrict__ vy,
                   uint32_t* __restrict__ sorted_id, Rec<Real>* __restrict__ rec) {
    const uint32_t s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= n) return;
    const uint32_t id = tmp_slot_id[s];
    const uint32_t c = cell_id[id];
    const uint32_t b = cell_start[c], e = cell_start[c + 1];
    uint32_t rank = 0;
    for (uint32_t t = b; t < e; ++t) rank += tmp_slot_id[t] < id ? 1u : 0u;   // ascending index inside the cell, src/core.zig:146-158
    const uint32_t dst = b + rank;
    sorted_id[dst] = id;
    st_rec(rec + dst, Rec<Real>{px[id], py[id], vx[id], vy[id]});
}

template <class Real>
__global__ void __launch_bounds__(256)
scatter_back_kernel(const uint32_t* __restrict__ sorted_id, const Rec<Real>* __restrict__ rec, uint32_t n,
                    Real* __restrict__ px, Real* __restrict__ py, Real* __restrict__ vx, Real* __restrict__ vy) {
    const uint32_t s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= n) return;
    const uint32_t id = sorted_id[s];
    const Rec<Real> r = ld_rec_cg(rec + s);
    px[id] = r.x; py[id] = r.y; vx[id] = r.u; vy[id] = r.w;
}

// ---- detect ----------------------------------------------------------------------------------
// Walk the contacts of slot `a` (cell column gx, row gy) in the order the reference's pair list
// mentions this particle.  Global list position = (centre slot, offset index o, neighbour slot)
// with offsets {(0,0),(0,1),(1,0),(1,1),(-1,1)} (src/core.zig:58) and cells column-major (:120):
//   1. a as NEIGHBOUR of earlier centres, ascending centre slot:
//        cell (gx-1,gy-1) via o=3, (gx-1,gy) via o=2, (gx,gy-1) via o=1, own cell q<a via o=0
//   2. a as CENTRE: o=0 own cell (q != a), o=1 (gx,gy+1), o=2 (gx+1,gy), o=3 (gx+1,gy+1), o=4 (gx-1,gy+1)
//   3. a as NEIGHBOUR of later centres: own cell q>a via o=0, cell (gx+1,gy-1) via o=4
// f(q, o, centre_flag) is called for every overlapping partner slot q (checkCollision, :339-346).
template <class Real, class F>
__device__ __forceinline__ void walk_contacts(uint32_t a, const Rec<Real>* __restrict__ rec, const uint32_t* __restrict__ cell_start,
                                              uint32_t c, const StepConst<Real>& k, F&& f) {
    const uint32_t rows = k.rows, cols = k.cols;
    const uint32_t gx = c / rows, gy = c - gx * rows;
    const auto pa = ld_pos(rec + a);
    const Real r2sq = k.r2sq;
    auto seg = [&](uint32_t b, uint32_t e, uint32_t o, uint32_t centre) {
        for (uint32_t q = b; q < e; ++q) {
            const auto pq = ld_pos(rec + q);
            // distanceSquared(pos_i, pos_j): (xj-xi)*(xj-xi) + (yj-yi)*(yj-yi); i = neighbour, j = centre  (:333-338)
            const Real dx = centre ? pa.x - pq.x : pq.x - pa.x;
            const Real dy = centre ? pa.y - pq.y : pq.y - pa.y;
            const Real d2 = dx * dx + dy * dy;
            if (d2 < r2sq) f(q, o, centre);
        }
    };
    const bool up = gy > 0, down = gy + 1 < rows, left = gx > 0, right = gx + 1 < cols;
    const uint32_t own_b = cell_start[c], own_e = cell_start[c + 1];
    // 1. neighbour of earlier centres
    if (left) {
        const uint32_t cl = c - rows;
        if (up) seg(cell_start[cl - 1], cell_start[cl], 3u, 0u);
        seg(cell_start[cl], cell_start[cl + 1], 2u, 0u);
    }
    if (up) seg(cell_start[c - 1], own_b, 1u, 0u);
    seg(own_b, a, 0u, 0u);
    // 2. centre
    seg(own_b, a, 0u, 1u);
    seg(a + 1, own_e, 0u, 1u);
    if (down) seg(own_e, cell_start[c + 2], 1u, 1u);
    if (right) {
        const uint32_t cr = c + rows;
        seg(cell_start[cr], cell_start[cr + 1], 2u, 1u);
        if (down) seg(cell_start[cr + 1], cell_start[cr + 2], 3u, 1u);
    }
    if (left && down) { const uint32_t cl = c - rows; seg(cell_start[cl + 1], cell_start[cl + 2], 4u, 1u); }
    // 3. neighbour of later centres
    seg(a + 1, own_e, 0u, 0u);
    if (right && up) { const uint32_t cr = c + rows; seg(cell_start[cr - 1], cell_start[cr], 4u, 0u); }
}

template <class Real>
__global__ void __launch_bounds__(256)
count_contacts_kernel(const Rec<Real>* __restrict__ rec, const uint32_t* __restrict__ cell_start,
                      const uint32_t* __restrict__ sorted_id, const uint32_t* __restrict__ cell_id, uint32_t n,
                      StepConst<Real> k, uint32_t* __restrict__ inc_count, uint32_t* __restrict__ ctr_count) {
    const uint32_t a = blockIdx.x * blockDim.x + threadIdx.x;
    if (a >= n) return;
    const uint32_t c = cell_id[sorted_id[a]];
    uint32_t ninc = 0, nctr = 0;
    walk_contacts<Real>(a, rec, cell_start, c, k, [&](uint32_t, uint32_t, uint32_t centre) { ++ninc; nctr += centre; });
    inc_count[a] = ninc;
    if (ctr_count) ctr_count[a] = nctr;
}

template <class Real>
__global__ void __launch_bounds__(256)
fill_contacts_kernel(const Rec<Real>* __restrict__ rec, const uint32_t* __restrict__ cell_start,
                     const uint32_t* __restrict__ sorted_id, const uint32_t* __restrict__ cell_id, uint32_t n,
                     StepConst<Real> k, const uint32_t* __restrict__ inc_off, uint32_t* __restrict__ entries, uint32_t entries_cap,
                     unsigned long long* __restrict__ head,
                     const uint32_t* __restrict__ ctr_off, uint64_t* __restrict__ pairs, uint32_t pairs_cap,
                     DevState* __restrict__ ds) {
    const uint32_t a = blockIdx.x * blockDim.x + threadIdx.x;
    if (a >= n) return;
    if (ds->inc_total > entries_cap) {                       // list does not fit: flag, leave the sub-step at its post-bounce state
        if (a == 0) atomicOr(&ds->flags, FLAG_CAPACITY);
        head[a] = (unsigned long long)inc_off[a + 1];        // empty list
        return;
    }
    const uint32_t id_a = sorted_id[a];
    const uint32_t c = cell_id[id_a];
    uint32_t w = inc_off[a];
    head[a] = (unsigned long long)w;                         // high word: 1 + round of the last advance (0 = none yet)
    const bool cap_pairs = pairs != nullptr && ds->ctr_total <= pairs_cap;
    uint32_t pw = cap_pairs ? ctr_off[a] : 0u;
    walk_contacts<Real>(a, rec, cell_start, c, k, [&](uint32_t q, uint32_t o, uint32_t centre) {
        entries[w++] = make_entry(q, o, centre);
        if (cap_pairs && centre) {                           // collision_pairs.append({particle_i, particle_j})  src/core.zig:203
            pairs[2 * (uint64_t)pw] = sorted_id[q];
            pairs[2 * (uint64_t)pw + 1] = id_a;
            ++pw;
        }
    });
}

// ---- resolve ---------------------------------------------------------------------------------
// resolve_collision(particles, i, j, dt, t_acc, particle_damping), src/core.zig:348-392.
template <class Real>
__device__ __forceinline__ void resolve_pair(Rec<Real>& I, Rec<Real>& J, const StepConst<Real>& k) {
    const Real ddx = J.x - I.x, ddy = J.y - I.y;
    const Real dist = sqrt_rn(ddx * ddx + ddy * ddy);                         // :354
    if (dist < (Real)0.0000000001) return;                                    // :355
    const Real nx = div_rn(ddx, dist), ny = div_rn(ddy, dist);                // :361
    const Real rvx = J.u - I.u, rvy = J.w - I.w;                              // :363
    const Real van = rvx * nx + rvy * ny;                                     // :365
    const Real kk = div_rn(k.k_coef * van, (Real)2);                          // :368
    const Real overlap = k.min_dis - dist;                                    // :372-374
    const Real corr = overlap * (Real)0.5;                                    // :375
    if (overlap > (Real)0) {                                                  // :377
        if (!k.t_acc) {
            I.x = I.x - corr * nx; I.y = I.y - corr * ny;                     // :379
            J.x = J.x + corr * nx; J.y = J.y + corr * ny;                     // :380
        } else {
            const Real pax = ((Real)10000.0 * nx) * overlap;                  // :385
            const Real pay = ((Real)10000.0 * ny) * overlap;
            I.u = I.u - pax * k.dt; I.w = I.w - pay * k.dt;                   // :386
            J.u = J.u + pax * k.dt; J.w = J.w + pay * k.dt;                   // :387
        }
    }
    I.u = I.u - nx * kk; I.w = I.w - ny * kk;                                 // :390
    J.u = J.u + nx * kk; J.w = J.w + ny * kk;                                 // :391
}

__device__ __forceinline__ void worklist_push(uint32_t* __restrict__ list, uint32_t* __restrict__ count, uint32_t v) {
    // warp-aggregated append
    const unsigned m = __activemask();
    const unsigned lane = threadIdx.x & 31;
    const int leader = __ffs(m) - 1;
    uint32_t base = 0;
    if ((int)lane == leader) base = atomicAdd(count, (uint32_t)__popc(m));
    base = __shfl_sync(m, base, leader);
    list[base + __popc(m & ((1u << lane) - 1))] = v;
}

template <class Real>
__global__ void __launch_bounds__(256)
resolve_kernel(Rec<Real>* __restrict__ rec, const uint32_t* __restrict__ inc_off, const uint32_t* __restrict__ entries,
               unsigned long long* head, uint32_t* wl0, uint32_t* wl1, uint32_t* wl2,
               uint32_t n, StepConst<Real> k, DevState* ds) {
    cg::grid_group grid = cg::this_grid();
    const uint32_t tid = blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t nthreads = gridDim.x * blockDim.x;
    uint32_t* lists[3] = {wl0, wl1, wl2};
    if (ds->flags & FLAG_CAPACITY) return;                   // uniform across the grid: set before this launch
    uint32_t r = 0;
    for (;; ++r) {
        const uint32_t cur = r % 3, nxt = (r + 1) % 3, nn = (r + 2) % 3;
        const uint32_t cnt = r == 0 ? n : __ldcg(&ds->wl_count[cur]);
        if (cnt == 0) break;
        if (tid == 0) ds->wl_count[nn] = 0;                  // consumed in round r-1, refilled in round r+1
        const uint32_t* wl = lists[cur];
        for (uint32_t idx = tid; idx < cnt; idx += nthreads) {
            const uint32_t a = r == 0 ? idx : __ldcg(wl + idx);
            const unsigned long long ha = __ldcg(head + a);
            if ((uint32_t)(ha >> 32) == r + 1) continue;     // a already ran a pair this round (executed by its partner)
            const uint32_t ia = (uint32_t)ha;
            const uint32_t end_a = inc_off[a + 1];
            if (ia >= end_a) continue;                       // a's list is finished
            const uint32_t ea = entries[ia];
            const uint32_t b = ea & ENT_SLOT_MASK;
            const unsigned long long hb = __ldcg(head + b);
            if ((uint32_t)(hb >> 32) == r + 1) continue;     // b moved this round; b is on the next list and will look at a
            const uint32_t ib = (uint32_t)hb;
            const uint32_t end_b = inc_off[b + 1];
            if (ib >= end_b) continue;
            const uint32_t eb = entries[ib];
            // same pair seen from the other side: partner = a, same offset, opposite role
            if (eb != (a | (ea & (7u << 28)) | (~ea & (1u << 31)))) continue;
            // Both particles on this round's list (b's head last moved in round r-1, or never when r = 0): the
            // lower slot runs the pair.  Membership is read from the SAME 64-bit word as the head index, so a
            // concurrent executor cannot make the two threads disagree.
            if (b < a && (uint32_t)(hb >> 32) == r) continue;
            const bool a_centre = (ea >> 31) != 0;
            const uint32_t si = a_centre ? b : a, sj = a_centre ? a : b;   // entry = {particle_i = neighbour, particle_j = centre}
            Rec<Real> I = ld_rec_cg(rec + si), J = ld_rec_cg(rec + sj);
            resolve_pair<Real>(I, J, k);
            st_rec(rec + si, I);
            st_rec(rec + sj, J);
            const unsigned long long stamp = (unsigned long long)(r + 1) << 32;
            head[a] = stamp | (ia + 1);
            head[b] = stamp | (ib + 1);
            if (ia + 1 < end_a) worklist_push(lists[nxt], &ds->wl_count[nxt], a);
            if (ib + 1 < end_b) worklist_push(lists[nxt], &ds->wl_count[nxt], b);
        }
        grid.sync();
    }
    if (tid == 0) {
        ds->rounds_last = r;
        ds->rounds_total += r;
        ds->pairs_total += ds->inc_total / 2;
        ds->wl_count[0] = ds->wl_count[1] = ds->wl_count[2] = 0;
    }
}

// ---- host side ---------------------------------------------------------------------------------
template <class Real>
struct Collider {
    uint64_t n = 0;
    uint32_t rows = 0, cols = 0, ncell = 0;
    bool capture = false;
    int device = 0;
    int resolve_grid = 0;

    uint32_t *cell_count = nullptr, *cell_start = nullptr, *scan_sums = nullptr;
    uint32_t *cell_id = nullptr, *arrival = nullptr, *tmp_slot_id = nullptr, *sorted_id = nullptr;
    Rec<Real>* rec = nullptr;
    uint32_t *inc_count = nullptr, *inc_off = nullptr, *ctr_count = nullptr, *ctr_off = nullptr;
    uint32_t* entries = nullptr; uint32_t entries_cap = 0;
    unsigned long long* head = nullptr;
    uint32_t* wl[3] = {nullptr, nullptr, nullptr};
    uint64_t* pairs = nullptr; uint32_t pairs_cap = 0;
    DevState* ds = nullptr;

    template <class T> int alloc(T** p, uint64_t count, Error& err) {
        PS_CUDA_TRY(err, cudaMalloc((void**)p, (count ? count : 1) * sizeof(T)));
        return PS_OK;
    }

    int create(const ps_config& cfg, uint64_t n_, uint32_t rows_, uint32_t cols_, cudaStream_t st, Error& err) {
        n = n_; rows = rows_; cols = cols_; ncell = rows * cols; device = cfg.device;
        const uint64_t per = cfg.contacts_per_particle ? cfg.contacts_per_particle : 12;
        uint64_t cap = per * (n ? n : 1);
        if (cap > 0xfffffff0ull) cap = 0xfffffff0ull;
        entries_cap = (uint32_t)cap;
        pairs_cap = entries_cap / 2;
        const uint64_t scan_len = (uint64_t)ncell > n ? ncell : n;
        int rc;
        if ((rc = alloc(&cell_count, (uint64_t)ncell, err))) return rc;
        if ((rc = alloc(&cell_start, (uint64_t)ncell + 3, err))) return rc;
        if ((rc = alloc(&scan_sums, scan_tiles(scan_len) + 1, err))) return rc;
        if ((rc = alloc(&cell_id, n, err))) return rc;
        if ((rc = alloc(&arrival, n, err))) return rc;
        if ((rc = alloc(&tmp_slot_id, n, err))) return rc;
        if ((rc = alloc(&sorted_id, n, err))) return rc;
        if ((rc = alloc(&rec, n, err))) return rc;
        if ((rc = alloc(&inc_count, n, err))) return rc;
        if ((rc = alloc(&inc_off, n + 1, err))) return rc;
        if ((rc = alloc(&entries, (uint64_t)entries_cap, err))) return rc;
        if ((rc = alloc(&head, n, err))) return rc;
        for (int i = 0; i < 3; ++i) if ((rc = alloc(&wl[i], n, err))) return rc;
        if ((rc = alloc(&ds, 1, err))) return rc;
        PS_CUDA_TRY(err, cudaMemsetAsync(ds, 0, sizeof(DevState), st));
        int per_sm = 0, sms = 0;
        PS_CUDA_TRY(err, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, resolve_kernel<Real>, 256, 0));
        PS_CUDA_TRY(err, cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device));
        int coop = 0;
        PS_CUDA_TRY(err, cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, device));
        if (!coop || per_sm < 1) { err.status = PS_ERR_CUDA; err.text = "device cannot run cooperative launches"; return err.status; }
        resolve_grid = per_sm * sms;
        return PS_OK;
    }

    void destroy() {
        cudaFree(cell_count); cudaFree(cell_start); cudaFree(scan_sums); cudaFree(cell_id); cudaFree(arrival);
        cudaFree(tmp_slot_id); cudaFree(sorted_id); cudaFree(rec); cudaFree(inc_count); cudaFree(inc_off);
        cudaFree(ctr_count); cudaFree(ctr_off); cudaFree(entries); cudaFree(head);
        for (auto& w : wl) cudaFree(w);
        cudaFree(pairs); cudaFree(ds);
    }

    int ensure_capture(Error& err) {
        if (pairs) return PS_OK;
        int rc;
        if ((rc = alloc(&ctr_count, n, err))) return rc;
        if ((rc = alloc(&ctr_off, n + 1, err))) return rc;
        if ((rc = alloc(&pairs, 2 * (uint64_t)pairs_cap, err))) return rc;
        return PS_OK;
    }

    // One iteration of the loop body at src/core.zig:247-267.
    int substep(Real* px, Real* py, Real* vx, Real* vy, const StepConst<Real>& k, cudaStream_t st, Error& err, uint64_t* launches) {
        if (n == 0) return PS_OK;
        if (capture) { const int rc = ensure_capture(err); if (rc) return rc; }
        const uint32_t n32 = (uint32_t)n;
        const unsigned gn = (unsigned)((n + 255) / 256);
        // shg.clear() (:262) — only the counters need it; every other array is fully rewritten
        PS_CUDA_TRY(err, cudaMemsetAsync(cell_count, 0, (uint64_t)ncell * sizeof(uint32_t), st));
        integrate_bin_kernel<Real><<<gn, 256, 0, st>>>(px, py, vx, vy, n32, k, cell_count, cell_id, arrival, ds);
        PS_CUDA_TRY(err, exclusive_scan_u32(cell_count, cell_start, ncell, scan_sums, cell_start + ncell, &ds->max_cell, st));
        place_kernel<<<gn, 256, 0, st>>>(cell_id, arrival, cell_start, n32, tmp_slot_id);
        rank_gather_kernel<Real><<<gn, 256, 0, st>>>(tmp_slot_id, cell_id, cell_start, n32, px, py, vx, vy, sorted_id, rec);
        count_contacts_kernel<Real><<<gn, 256, 0, st>>>(rec, cell_start, sorted_id, cell_id, n32, k, inc_count, capture ? ctr_count : nullptr);
        PS_CUDA_TRY(err, exclusive_scan_u32(inc_count, inc_off, n, scan_sums, &ds->inc_total, nullptr, st));
        PS_CUDA_TRY(err, cudaMemcpyAsync(inc_off + n, &ds->inc_total, sizeof(uint32_t), cudaMemcpyDeviceToDevice, st));
        if (capture) {
            PS_CUDA_TRY(err, exclusive_scan_u32(ctr_count, ctr_off, n, scan_sums, &ds->ctr_total, nullptr, st));
        }
        fill_contacts_kernel<Real><<<gn, 256, 0, st>>>(rec, cell_start, sorted_id, cell_id, n32, k, inc_off, entries, entries_cap,
                                                       head, capture ? ctr_off : nullptr, capture ? pairs : nullptr, pairs_cap, ds);
        {
            Rec<Real>* a_rec = rec; const uint32_t* a_off = inc_off; const uint32_t* a_ent = entries;
            unsigned long long* a_head = head; uint32_t *a0 = wl[0], *a1 = wl[1], *a2 = wl[2];
            uint32_t a_n = n32; StepConst<Real> a_k = k; DevState* a_ds = ds;
            void* args[] = {&a_rec, &a_off, &a_ent, &a_head, &a0, &a1, &a2, &a_n, &a_k, &a_ds};
            const uint64_t want = (n + 255) / 256;
            const unsigned grid = (unsigned)(want < (uint64_t)resolve_grid ? want : (uint64_t)resolve_grid);
            PS_CUDA_TRY(err, cudaLaunchCooperativeKernel((const void*)resolve_kernel<Real>, dim3(grid), dim3(256), args, 0, st));
        }
        scatter_back_kernel<Real><<<gn, 256, 0, st>>>(sorted_id, rec, n32, px, py, vx, vy);
        PS_CUDA_TRY(err, cudaGetLastError());
        *launches += 13 + (capture ? 3 : 0);   // bin, 3+3 scan, place, rank, count, fill, resolve, scatter-back (+3 scan when capturing)
        return PS_OK;
    }

    int check_flags(cudaStream_t st, Error& err) {
        DevState h{};
        PS_CUDA_TRY(err, cudaMemcpyAsync(&h, ds, sizeof(DevState), cudaMemcpyDeviceToHost, st));
        PS_CUDA_TRY(err, cudaStreamSynchronize(st));
        if (h.flags & FLAG_CAPACITY) {
            err.status = PS_ERR_CAPACITY;
            err.text = "contact list exceeded its capacity (" + std::to_string(entries_cap) + " entries); raise ps_config.contacts_per_particle";
            return err.status;
        }
        if (h.flags & FLAG_CELL_OVERFLOW) {
            err.status = PS_ERR_CELL_OVERFLOW;
            err.text = "a grid cell held more than 255 particles; the reference's u8 counter (src/core.zig:51,132) cannot represent this scene";
            return err.status;
        }
        return PS_OK;
    }

    int read_stats(ps_stats* out, cudaStream_t st, Error& err) {
        DevState h{};
        PS_CUDA_TRY(err, cudaMemcpyAsync(&h, ds, sizeof(DevState), cudaMemcpyDeviceToHost, st));
        PS_CUDA_TRY(err, cudaStreamSynchronize(st));
        out->pairs_last = h.inc_total / 2;
        out->pairs_total = h.pairs_total;
        out->rounds_last = h.rounds_last;
        out->rounds_total = h.rounds_total;
        out->max_cell_count = h.max_cell;
        return PS_OK;
    }

    int read_pairs(uint64_t* out, uint64_t cap, uint64_t* count, cudaStream_t st, Error& err) {
        if (capture && n == 0) { *count = 0; return PS_OK; }   // an empty scene has an empty list
        if (!capture || !pairs) { err.status = PS_ERR_BAD_ARG; err.text = "ps_contact_pairs: call ps_capture_pairs(sim, 1) before stepping"; return err.status; }
        DevState h{};
        PS_CUDA_TRY(err, cudaMemcpyAsync(&h, ds, sizeof(DevState), cudaMemcpyDeviceToHost, st));
        PS_CUDA_TRY(err, cudaStreamSynchronize(st));
        *count = h.ctr_total;
        if (h.ctr_total > pairs_cap) { err.status = PS_ERR_CAPACITY; err.text = "pair capture buffer too small"; return err.status; }
        const uint64_t m = cap < h.ctr_total ? cap : h.ctr_total;
        if (m) {
            PS_CUDA_TRY(err, cudaMemcpyAsync(out, pairs, m * 2 * sizeof(uint64_t), cudaMemcpyDeviceToHost, st));
            PS_CUDA_TRY(err, cudaStreamSynchronize(st));
        }
        return PS_OK;
    }
};

}  // namespace ps
