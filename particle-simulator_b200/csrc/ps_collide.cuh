// ps_collide.cuh — the collision half of updatePhysics (src/core.zig:262-266): uniform-grid
// broadphase, ordered contact detection and the exact-order resolve.
//
// The reference resolves an ordered pair list sequentially (Gauss-Seidel, src/core.zig:211-213);
// the result depends on that order, so this file reproduces the order instead of
// approximating it.
//
// State layout.  While a collision scene is being stepped the particles live in `rec`: one
// 16/32-byte record {px,py,vx,vy} per particle, kept PHYSICALLY IN CELL ORDER ("slots") from one
// sub-step to the next, with `pid[slot]` = original particle index.  Particles move less than half
// a cell per sub-step, so the previous order is almost the next one: every pass below reads and
// writes memory nearly sequentially, and the atomics of the counting sort hit neighbouring
// counters.  The SoA arrays in original index order are rebuilt only when the caller asks for the
// state (download / ps_device_state), SURVEY Q9.
//
//   bin      integrate + bounce in registers (ps_integrate.cuh arithmetic), computeCoordinate /
//            computeGridIndex (:103-121), per-cell count (:124-134, one atomic per particle)
//   scan     computeStartArray (:137-142), ps_scan.cuh
//   place    particle indices into their cell's slot range, unordered
//   move     computeSortedArray (:145-159): rank each particle among its cell's indices (ascending
//            index = what the reference's stable counting sort produces) and write its integrated
//            record and index to slot cell_start + rank of the other buffer
//   detect   every slot tests the candidates of its 3x3 cell block once; the hits, as a bit mask, are permuted into the
//            order in which the reference's pair list (:164-209) mentions that particle.  The position of an entry in
//            the global list is the lexicographic key (centre slot, offset index, neighbour slot), so each particle's
//            own sub-list is produced without materialising or sorting the global list.  Entry 0 goes into the
//            particle's 64-bit head word, entries 1-8 into its own chunk at a fixed place: nothing is reserved or
//            scanned.  Warp-autonomous, two phases (all slots; then only the third that hit something, compacted).
//   resolve  a pair may run as soon as it heads BOTH of its particles' sub-lists: pairs that share
//            no particle commute, so executing all such "ready" pairs together, round after round,
//            is bit-identical to the sequential loop.  One persistent cooperative kernel runs the
//            rounds (one grid barrier each) over work lists that hold only the particles whose head
//            changed in the previous round, so total work is O(pairs).  Short lists are finished by
//            one CTA out of shared memory, one CTA barrier per round (deep chains = many short rounds).
#pragma once
#include <cooperative_groups.h>

#include <cstdio>
#include <cstdlib>

#include "ps_common.cuh"
#include "ps_integrate.cuh"
#include "ps_scan.cuh"

namespace ps {

namespace cg = cooperative_groups;

// ---- record types -------------------------------------------------------------------------
template <class Real> struct Rec;   // one particle: {px, py, vx, vy}
template <> struct __align__(16) Rec<float>  { float  x, y, u, w; };
template <> struct __align__(16) Rec<double> { double x, y, u, w; };

__device__ __forceinline__ Rec<float> ld_rec(const Rec<float>* p) {
    const float4 v = *reinterpret_cast<const float4*>(p);
    return Rec<float>{v.x, v.y, v.z, v.w};
}
__device__ __forceinline__ Rec<double> ld_rec(const Rec<double>* p) {
    const double2 a = reinterpret_cast<const double2*>(p)[0];
    const double2 b = reinterpret_cast<const double2*>(p)[1];
    return Rec<double>{a.x, a.y, b.x, b.y};
}
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
    uint32_t entries_used;     // contact-list entries written by detect in this sub-step (= 2 * pairs)
    uint32_t inc_total;        // entries_used of the last completed sub-step
    uint32_t ctr_total;        // pairs of this sub-step (pair capture only)
    uint32_t wl_count[3];      // rotating work-list lengths
    uint32_t rounds_last;
    uint32_t pool_used;        // words of the overflow pool handed out in this sub-step (sub-lists longer than 1 + LIST_CHUNK)
    unsigned long long pairs_total, rounds_total;
    uint32_t dbg[4];           // first offending slot / cell when FLAG_CORRUPT is raised
    uint32_t trace_on;         // PS_RESOLVE_TRACE: record the length and start time of every resolve round
    uint32_t trace_rounds;
    uint32_t trace_cnt[64];
    uint32_t trace_ns[64];     // nanoseconds since the kernel's first round
};
__device__ __forceinline__ unsigned long long global_ns() { unsigned long long t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)); return t; }
constexpr uint32_t FLAG_CELL_OVERFLOW = 1u, FLAG_CAPACITY = 2u, FLAG_CORRUPT = 4u, FLAG_LIST_TOO_LONG = 8u;

// Contact sub-lists.  Every particle's sub-list is the sequence of pair-list entries that mention it, in list order.
//   entry word   partner slot (bits 0..27) | role (bit 28: 1 = this particle is the centre) | LAST (bit 29: final entry of
//                the sub-list).  The offset index is not stored: a partner from another cell appears once in a sub-list, a
//                partner from the own cell twice with opposite roles, so partner + role identify the entry.
//   where        entry 0 lives in the particle's head word (below); entries 1..8 in the particle's own chunk of LIST_CHUNK
//                words at a fixed place (chunks[slot * 8 ..]), so most lists need no allocation, no prefix sum and no
//                cursor; a longer list keeps entries 1..7 in the chunk, the index of its overflow run in the pool in chunk
//                word 7, and entries 8.. in that run (one atomic per such particle).
constexpr uint32_t ENT_SLOT_MASK = (1u << 28) - 1;
constexpr int ENT_ROLE_BIT = 28;
constexpr uint32_t ENT_CENTRE = 1u << ENT_ROLE_BIT;
constexpr uint32_t ENT_LAST = 1u << 29;
constexpr uint32_t LIST_CHUNK = 8;
__device__ __forceinline__ uint32_t make_entry(uint32_t partner, uint32_t /*o*/, uint32_t centre) { return partner | (centre << ENT_ROLE_BIT); }

// x / grid_size (src/core.zig:104-105).  When grid_size is a power of two (the reference's 4.0) the quotient equals
// x * (1 / grid_size) bit for bit (scaling by a power of two is exact), which spares the IEEE division sequence.
template <class Real>
__device__ __forceinline__ Real cell_coord(Real x, const StepConst<Real>& k) {
    return k.gs_pow2 ? x * k.inv_gs : div_rn(x, k.grid_size);
}

// computeCoordinate (src/core.zig:103-117): column i and row j of the cell
template <class Real>
__device__ __forceinline__ void cell_xy(Real x, Real y, const StepConst<Real>& k, uint32_t& i, uint32_t& j) {
    const Real fi = floor_r(cell_coord<Real>(x, k));
    const Real fj = floor_r(cell_coord<Real>(y, k));
    // positions are inside [r, dim-r] here; the clamp mirrors @min(i, cols-1) and guards NaN / negatives
    i = fi > (Real)0 ? (fi < (Real)k.cols ? (uint32_t)fi : k.cols - 1) : 0u;
    j = fj > (Real)0 ? (fj < (Real)k.rows ? (uint32_t)fj : k.rows - 1) : 0u;
}

// computeCoordinate + computeGridIndex (src/core.zig:103-121)
template <class Real>
__device__ __forceinline__ uint32_t cell_of(Real x, Real y, const StepConst<Real>& k) {
    uint32_t i, j;
    cell_xy<Real>(x, y, k, i, j);
    return i * k.rows + j;
}

// the integrate loop + _applyBoundary on one record (src/core.zig:248-261)
template <class Real>
__device__ __forceinline__ void integrate_rec(Rec<Real>& r, const StepConst<Real>& k) {
    step_component<Real>(r.x, r.u, k.ax_dt, k.dt, k.lo, k.hi_x, k.neg_damp);
    step_component<Real>(r.y, r.w, k.ay_dt, k.dt, k.lo, k.hi_y, k.neg_damp);
}

// ---- SoA (original index order) <-> records ---------------------------------------------------
template <class Real>
__global__ void __launch_bounds__(256)
soa_to_rec_kernel(const Real* __restrict__ px, const Real* __restrict__ py, const Real* __restrict__ vx, const Real* __restrict__ vy,
                  uint32_t n, Rec<Real>* __restrict__ rec, uint32_t* __restrict__ pid) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    st_rec(rec + i, Rec<Real>{px[i], py[i], vx[i], vy[i]});
    pid[i] = i;
}

template <class Real>
__global__ void __launch_bounds__(256)
rec_to_soa_kernel(const uint32_t* __restrict__ pid, const Rec<Real>* __restrict__ rec, uint32_t n,
                  Real* __restrict__ px, Real* __restrict__ py, Real* __restrict__ vx, Real* __restrict__ vy) {
    const uint32_t s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= n) return;
    const uint32_t id = pid[s];
    const Rec<Real> r = ld_rec(rec + s);
    px[id] = r.x; py[id] = r.y; vx[id] = r.u; vy[id] = r.w;
}

// ---- bin: integrate + bounce + cell key + per-cell count ------------------------------------
constexpr int SORT_ILP = 2;
template <class Real>
__global__ void __launch_bounds__(256)
bin_kernel(const Rec<Real>* __restrict__ rec_old, uint32_t n, StepConst<Real> k, uint32_t* __restrict__ cell_count,
           uint32_t* __restrict__ slot_cell, uint32_t* __restrict__ arrival, DevState* __restrict__ ds) {
    // SORT_ILP slots per thread, CTA-strided: the loads, then the counting atomics, of all of them are in flight together
    const uint32_t s0 = blockIdx.x * (blockDim.x * SORT_ILP) + threadIdx.x;
    if (s0 == 0) { ds->entries_used = 0; ds->pool_used = 0; }   // detect of this sub-step counts and allocates from zero
    Rec<Real> r[SORT_ILP];
    uint32_t c[SORT_ILP], a[SORT_ILP];
#pragma unroll
    for (int j = 0; j < SORT_ILP; ++j) { const uint32_t s = s0 + j * blockDim.x; if (s < n) r[j] = ld_rec(rec_old + s); }
#pragma unroll
    for (int j = 0; j < SORT_ILP; ++j) {
        const uint32_t s = s0 + j * blockDim.x;
        if (s < n) {
            integrate_rec<Real>(r[j], k);
            c[j] = cell_of<Real>(r[j].x, r[j].y, k);
            a[j] = cell_count_fetch_inc(cell_count, c[j]);      // grid_array[c] += 1   src/core.zig:132 (u8 counters, :51)
        }
    }
#pragma unroll
    for (int j = 0; j < SORT_ILP; ++j) {
        const uint32_t s = s0 + j * blockDim.x;
        if (s < n) {
            slot_cell[s] = c[j];
            arrival[s] = a[j];
            if (a[j] >= 255u) atomicOr(&ds->flags, FLAG_CELL_OVERFLOW);   // the reference's u8 would wrap here
        }
    }
}

// ---- place + move: the stable counting sort ----------------------------------------------------
__global__ void __launch_bounds__(256)
place_kernel(const uint32_t* __restrict__ slot_cell, const uint32_t* __restrict__ arrival, const uint32_t* __restrict__ pid_old,
             const uint32_t* __restrict__ cell_start, uint32_t n, uint32_t* __restrict__ tmp_id) {
    const uint32_t s0 = blockIdx.x * (blockDim.x * SORT_ILP) + threadIdx.x;
    uint32_t c[SORT_ILP], a[SORT_ILP], id[SORT_ILP], b[SORT_ILP];
#pragma unroll
    for (int j = 0; j < SORT_ILP; ++j) {
        const uint32_t s = s0 + j * blockDim.x;
        if (s < n) { c[j] = slot_cell[s]; a[j] = arrival[s]; id[j] = pid_old[s]; }
    }
#pragma unroll
    for (int j = 0; j < SORT_ILP; ++j) if (s0 + j * blockDim.x < n) b[j] = cell_start[c[j]];
#pragma unroll
    for (int j = 0; j < SORT_ILP; ++j) {
        if (s0 + j * blockDim.x < n) {
            const uint32_t dst = b[j] + a[j];
            if (dst < n) tmp_id[dst] = id[j];                   // out of range only after a flagged counter overflow
        }
    }
}

template <class Real>
__global__ void __launch_bounds__(256)
move_kernel(const Rec<Real>* __restrict__ rec_old, const uint32_t* __restrict__ pid_old, const uint32_t* __restrict__ slot_cell,
            const uint32_t* __restrict__ cell_start, const uint32_t* __restrict__ tmp_id, uint32_t n, StepConst<Real> k,
            Rec<Real>* __restrict__ rec_new, uint32_t* __restrict__ pid_new, const DevState* __restrict__ ds) {
    const uint32_t s0 = blockIdx.x * (blockDim.x * SORT_ILP) + threadIdx.x;
    if (ds->flags) {                                         // an earlier sub-step failed: the state stays frozen where it failed
#pragma unroll
        for (int j = 0; j < SORT_ILP; ++j) {
            const uint32_t s = s0 + j * blockDim.x;
            if (s < n) { st_rec(rec_new + s, ld_rec(rec_old + s)); pid_new[s] = pid_old[s]; }
        }
        return;
    }
    uint32_t id[SORT_ILP], c[SORT_ILP], b[SORT_ILP], e[SORT_ILP];
    Rec<Real> r[SORT_ILP];
#pragma unroll
    for (int j = 0; j < SORT_ILP; ++j) {
        const uint32_t s = s0 + j * blockDim.x;
        if (s < n) { id[j] = pid_old[s]; c[j] = slot_cell[s]; r[j] = ld_rec(rec_old + s); }
    }
#pragma unroll
    for (int j = 0; j < SORT_ILP; ++j) if (s0 + j * blockDim.x < n) { b[j] = cell_start[c[j]]; e[j] = cell_start[c[j] + 1]; }
#pragma unroll
    for (int j = 0; j < SORT_ILP; ++j) {
        if (s0 + j * blockDim.x >= n) continue;
        uint32_t rank = 0;
        for (uint32_t t = b[j]; t < e[j] && t < n; ++t) rank += tmp_id[t] < id[j] ? 1u : 0u;   // ascending index inside the cell, src/core.zig:146-158
        integrate_rec<Real>(r[j], k);                                     // same arithmetic as bin_kernel, bit for bit
        if (b[j] + rank >= n) continue;                                   // only after a flagged counter overflow
        st_rec(rec_new + b[j] + rank, r[j]);
        pid_new[b[j] + rank] = id[j];
    }
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
struct Stencil {   // slot ranges of the 3x3 block around one cell, loaded once and reused by both walks
    uint32_t l0, l1, l2, l3;   // column gx-1: starts of rows gy-1, gy, gy+1, gy+2
    uint32_t m0, m1, m2, m3;   // column gx
    uint32_t r0, r1, r2, r3;   // column gx+1
    bool up, down, left, right;
};

template <class Real>
__device__ __forceinline__ Stencil load_stencil(const uint32_t* __restrict__ cell_start, uint32_t c, const StepConst<Real>& k);   // by cell index, below

template <class Real>
__device__ __forceinline__ Stencil load_stencil(const uint32_t* __restrict__ cell_start, uint32_t gx, uint32_t gy, uint32_t ncols,
                                                const StepConst<Real>& k) {
    Stencil s{};
    const uint32_t rows = k.rows, cols = ncols;
    const uint32_t c = gx * rows + gy;
    s.up = gy > 0; s.down = gy + 1 < rows; s.left = gx > 0; s.right = gx + 1 < cols;
    s.m1 = cell_start[c]; s.m2 = cell_start[c + 1];
    s.m0 = s.up ? cell_start[c - 1] : s.m1;
    s.m3 = s.down ? cell_start[c + 2] : s.m2;
    if (s.left) {
        const uint32_t cl = c - rows;
        s.l1 = cell_start[cl]; s.l2 = cell_start[cl + 1];
        s.l0 = s.up ? cell_start[cl - 1] : s.l1;
        s.l3 = s.down ? cell_start[cl + 2] : s.l2;
    }
    if (s.right) {
        const uint32_t cr = c + rows;
        s.r1 = cell_start[cr]; s.r2 = cell_start[cr + 1];
        s.r0 = s.up ? cell_start[cr - 1] : s.r1;
        s.r3 = s.down ? cell_start[cr + 2] : s.r2;
    }
    return s;
}

template <class Real>
__device__ __forceinline__ Stencil load_stencil(const uint32_t* __restrict__ cell_start, uint32_t c, const StepConst<Real>& k) {
    const uint32_t gx = c / k.rows;
    return load_stencil<Real>(cell_start, gx, c - gx * k.rows, k.cols, k);
}

template <class Real, class F>
__device__ __forceinline__ void walk_contacts(uint32_t a, const Rec<Real>* __restrict__ rec, const Stencil& s, Real r2sq, F&& f) {
    const auto pa = ld_pos(rec + a);
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
    // 1. neighbour of earlier centres (empty ranges when the cell does not exist)
    if (s.left) { seg(s.l0, s.l1, 3u, 0u); seg(s.l1, s.l2, 2u, 0u); }
    seg(s.m0, s.m1, 1u, 0u);
    seg(s.m1, a, 0u, 0u);
    // 2. centre
    seg(s.m1, a, 0u, 1u);
    seg(a + 1, s.m2, 0u, 1u);
    seg(s.m2, s.m3, 1u, 1u);
    if (s.right) { seg(s.r1, s.r2, 2u, 1u); seg(s.r2, s.r3, 3u, 1u); }
    if (s.left) seg(s.l2, s.l3, 4u, 1u);
    // 3. neighbour of later centres
    seg(a + 1, s.m2, 0u, 0u);
    if (s.right) seg(s.r0, s.r1, 4u, 0u);
}

// Per-particle head word, one 64-bit value so that a reader always sees a consistent snapshot:
//   bits 63..45  stamp = 1 + round of the last advance (0 = not advanced yet)
//   bits 44..32  pos   = index in the sub-list of the current head entry (<= 8191)
//   bit  31      LIVE  = the sub-list has a current head entry (0: empty, finished, or blocked in a slab's phase A)
//   bit  30      OVF   = the sub-list is longer than 1 + LIST_CHUNK entries (chunk word 7 is a pool index)
//   bits 29..0   the current head entry itself (partner, role, LAST)
// Whether a pair is ready is decided from the two head words alone; the thread that runs the pair fetches each particle's
// next entry from that particle's chunk (one 32-byte sector per particle, the same one every time).
constexpr int HEAD_POS_BITS = 13;
constexpr uint32_t HEAD_REM_MAX = (1u << HEAD_POS_BITS) - 1;            // longest sub-list
constexpr uint32_t HEAD_STAMP_MAX = (1u << (32 - HEAD_POS_BITS)) - 1;
constexpr uint32_t HEAD_LIVE = 1u << 31, HEAD_OVF = 1u << 30;
__host__ __device__ __forceinline__ unsigned long long head_word(uint32_t stamp, uint32_t pos, uint32_t low) {
    return ((unsigned long long)stamp << (32 + HEAD_POS_BITS)) | ((unsigned long long)pos << 32) | low;
}
__device__ __forceinline__ uint32_t head_stamp(unsigned long long h) { return (uint32_t)(h >> (32 + HEAD_POS_BITS)); }
__device__ __forceinline__ uint32_t head_pos(unsigned long long h) { return (uint32_t)(h >> 32) & HEAD_REM_MAX; }
__device__ __forceinline__ uint32_t head_low(unsigned long long h) { return (uint32_t)h; }
// entry j >= 1 of slot a's sub-list
__device__ __forceinline__ uint32_t load_entry(const uint32_t* __restrict__ chunks, const uint32_t* __restrict__ pool, uint32_t a,
                                               uint32_t j, uint32_t ovf) {
    const uint32_t idx = j - 1u;
    const uint32_t* c = chunks + (size_t)a * LIST_CHUNK;
    if (!ovf || idx < LIST_CHUNK - 1u) return c[idx];
    return pool[c[LIST_CHUNK - 1u] + (idx - (LIST_CHUNK - 1u))];
}

// detect_kernel builds every particle's contact sub-list.  A warp owns a span of consecutive slots and works alone
// (no CTA barrier, no CTA-wide scan):
//   phase 1  every slot of the span, 32 at a time: cell, the 3x3 block's slot ranges, every candidate tested ONCE in
//            one flat loop over the three columns' ranges; slots with at least one hit (about a third of them at the
//            reference density) are queued in shared memory as {slot, hit mask}
//   phase 2  the queue, 32 at a time, so every lane has work: the hit mask is permuted into the reference's list order,
//            entry 0 goes into the head word, the others into the slot's own chunk (nothing is reserved or scanned);
//            particles that go on the first work list are compacted in place in the queue
//   finally  one atomic reserves the span's piece of the first work list, written as one coalesced run.
// A particle whose 3x3 block holds more than 32 candidates takes a generic two-walk path.
//
// Multi-GPU slabs (SLAB = true, ps_slab.cuh): the local grid has one ghost column on each side that
// has a neighbour rank.  An entry is listed on this rank only if its CENTRE particle sits in an own
// column (the pair list is ordered by centre, so a pair belongs to the rank that owns its centre);
// particles of the first two local columns start "blocked" (their sub-list becomes visible only after
// the left rank has handed over their state): their head word is written without LIVE and `brem` is set.
struct SlabView {
    uint32_t gl, gr;          // 1 if a left / right ghost column exists
    uint32_t ncl;             // local columns (own + ghosts)
    uint32_t col_off;         // global column of local column 0
    uint32_t gcols;           // global column count (clamp of computeCoordinate, src/core.zig:108)
    const uint32_t* n_local;  // device: slots in use
    const uint32_t* n_blocked;// device: slots of the first two local columns (0 when gl = 0)
    uint32_t* brem;           // per slot: sub-list length of a blocked particle
};

struct OwnFlags { bool a_own, l_own, r_own; };   // which columns of the 3x3 block belong to this rank (all true on one GPU)

#ifndef PS_DETECT_UNROLL
#define PS_DETECT_UNROLL 2      // tuning knobs of detect_kernel.  Measured on B200 at 16 Mi particles, whole sub-step: unroll 1 / 4: +1.6 %;
                                // 5 or 8 CTAs per SM (48 / 32 registers): +4 % / -1 %; 128- or 512-thread CTAs: +4 % / +3 %
#endif
#ifndef PS_DETECT_MINB
#define PS_DETECT_MINB 8      // CTAs of 256 threads per SM (fp32): the kernel waits on loads, so residency beats registers (measured at 16 Mi
                              // particles, whole sub-step: 5 -> 1.147 ms, 6 -> 1.117 ms, 8 -> 1.106 ms)
#endif
#ifndef PS_DETECT_THREADS
#define PS_DETECT_THREADS 256
#endif
constexpr int kDetectUnroll = PS_DETECT_UNROLL;
constexpr int DETECT_THREADS = PS_DETECT_THREADS;   // CTA size of detect_kernel: the unit of list-space reservation (one scan + two atomics per CTA)
// ---- the usual case: at most 31 candidates ------------------------------------------------------------------
// One flat loop tests the candidates of all three columns (iterations per warp = the largest block among its lanes,
// not the sum of the largest columns); bit t of the result = t-th candidate in column order (left, own, right).
template <class Real>
__device__ __forceinline__ uint32_t test_flat(const Rec<Real>* __restrict__ rec, uint32_t a, decltype(ld_pos((const Rec<Real>*)nullptr)) pa,
                                              const Stencil& st, uint32_t nL, uint32_t nM, uint32_t total, Real r2sq) {
    const uint32_t bL = st.l0, bM = st.m0 - nL, bR = st.r0 - nL - nM, nLM = nL + nM;   // wrap-around arithmetic: slot = base + t
    uint32_t hits = 0;
#pragma unroll kDetectUnroll
    for (uint32_t t = 0; t < total; ++t) {
        const uint32_t q = t + (t < nL ? bL : (t < nLM ? bM : bR));
        const auto pq = ld_pos(rec + q);
        const Real dx = pq.x - pa.x, dy = pq.y - pa.y;      // (xj-xi)^2 does not depend on which of the two is i (:333-338)
        const Real d2 = dx * dx + dy * dy;
        hits |= (d2 < r2sq ? 1u : 0u) << t;
    }
    return hits & ~(1u << (nL + (a - st.m0)));              // pi != pj (:199)
}

// The sub-list of slot `a` as eight runs of consecutive slots, in list order (the eleven segments documented at
// walk_contacts; neighbouring segments that are contiguous in slot space and share the role are one run):
//   0 [l0,l2) neighbour   1 [m0,a) neighbour   2 [m1,a) centre    3 (a,m3) centre
//   4 [r1,r3) centre      5 [l2,l3) centre     6 (a,m2) neighbour 7 [r0,r1) neighbour
// ListRuns holds where each run starts in "list space" (bit position in the list-ordered mask).
struct ListRuns {
    uint32_t t1, t2, t3, t4, t5, t6, t7, t8;               // start of runs 1..7 and the total length (run 0 starts at 0)
};
__device__ __forceinline__ ListRuns list_runs(const Stencil& st, uint32_t a) {
    ListRuns r;
    r.t1 = st.l2 - st.l0;
    r.t2 = r.t1 + (a - st.m0);
    r.t3 = r.t2 + (a - st.m1);
    r.t4 = r.t3 + (st.m3 - a - 1u);
    r.t5 = r.t4 + (st.r3 - st.r1);
    r.t6 = r.t5 + (st.l3 - st.l2);
    r.t7 = r.t6 + (st.m2 - a - 1u);
    r.t8 = r.t7 + (st.r1 - st.r0);
    return r;
}

// Reorder the column-ordered hit mask into list order (own-cell partners appear twice, Q1).  All shifts are < 32 and all
// lengths <= 31 because the caller takes this path only when r.t8 <= 31.  `own` drops the runs whose centre lies in a
// column of another rank (multi-GPU slabs).
__device__ __forceinline__ uint32_t list_mask(uint32_t hits, const Stencil& st, uint32_t a, uint32_t nL, uint32_t nM,
                                              const ListRuns& r, OwnFlags own) {
    auto run = [&](uint32_t src, uint32_t len, uint32_t dst) { return ((hits >> src) & ((1u << len) - 1u)) << dst; };
    const uint32_t am = nL + (a - st.m0);                   // bit of `a` itself in column order
    uint32_t left = run(0u, r.t1, 0u);
    uint32_t mid = run(nL, r.t2 - r.t1, r.t1) | run(nL + (st.m1 - st.m0), r.t3 - r.t2, r.t2) | run(am + 1u, r.t4 - r.t3, r.t3) |
                   run(nL + nM + (st.r1 - st.r0), r.t5 - r.t4, r.t4) | run(r.t1, r.t6 - r.t5, r.t5) | run(am + 1u, r.t7 - r.t6, r.t6);
    uint32_t right = run(nL + nM, r.t8 - r.t7, r.t7);
    if (!own.l_own) left = 0u;
    if (!own.a_own) mid = 0u;
    if (!own.r_own) right = 0u;
    return left | mid | right;
}

// Where entry j >= 1 of a sub-list of `cnt` entries goes: the slot's chunk, or the overflow run at `run` in the pool.
__device__ __forceinline__ void store_entry(uint32_t* __restrict__ chunk, uint32_t* __restrict__ pool, bool ovf, uint32_t run, uint32_t j, uint32_t e) {
    const uint32_t idx = j - 1u;
    if (!ovf || idx < LIST_CHUNK - 1u) chunk[idx] = e;
    else pool[run + (idx - (LIST_CHUNK - 1u))] = e;
}

// Write the entries of a list-ordered mask (entry 0 is returned, for the head word; the last one carries ENT_LAST).
// Entry of list bit p in run i = base_i + p with base_i = (first slot of the run - start of the run) + role bit, all
// modulo 2^32.
__device__ __forceinline__ uint32_t emit_list(uint32_t lm, const Stencil& st, uint32_t a, const ListRuns& r,
                                              uint32_t* __restrict__ chunk, uint32_t* __restrict__ pool, bool ovf, uint32_t run) {
    const uint32_t b0 = st.l0, b1 = st.m0 - r.t1, b2 = st.m1 - r.t2 + ENT_CENTRE, b3 = a + 1u - r.t3 + ENT_CENTRE,
                   b4 = st.r1 - r.t4 + ENT_CENTRE, b5 = st.l2 - r.t5 + ENT_CENTRE, b6 = a + 1u - r.t6, b7 = st.r0 - r.t7;
    uint32_t e0 = 0u, j = 0u;
    while (lm) {
        const uint32_t p = (uint32_t)__ffs((int)lm) - 1u;
        lm &= lm - 1u;
        const uint32_t lo = p < r.t2 ? (p < r.t1 ? b0 : b1) : (p < r.t3 ? b2 : b3);
        const uint32_t hi = p < r.t6 ? (p < r.t5 ? b4 : b5) : (p < r.t7 ? b6 : b7);
        const uint32_t e = ((p < r.t4 ? lo : hi) + p) | (lm ? 0u : ENT_LAST);
        if (j == 0u) e0 = e; else store_entry(chunk, pool, ovf, run, j, e);
        ++j;
    }
    return e0;
}

constexpr uint32_t DETECT_SPAN_MAX = 256;   // slots per warp (a multiple of 32); the launch picks a shorter span for small scenes
constexpr uint32_t DETECT_SLOW = 1u << 31;  // queue item flag: more than 32 candidates, take the generic walk

// The cell of slot a's particle and the slot ranges of its 3x3 block (both phases need them; phase 2 recomputes them
// for the third of the slots that got that far instead of carrying twelve words per slot through shared memory).
template <class Real, bool SLAB>
__device__ __forceinline__ Stencil locate(decltype(ld_pos((const Rec<Real>*)nullptr)) pa, const uint32_t* __restrict__ cell_start,
                                          const StepConst<Real>& k, const SlabView& sv, DevState* __restrict__ ds, uint32_t a, uint32_t n,
                                          OwnFlags& own) {
    if (SLAB) {
        // local cell: global computeCoordinate, then shift by the slab's first local column
        const Real fi = floor_r(cell_coord<Real>(pa.x, k)), fj = floor_r(cell_coord<Real>(pa.y, k));
        const uint32_t gi = fi > (Real)0 ? (fi < (Real)sv.gcols ? (uint32_t)fi : sv.gcols - 1) : 0u;
        uint32_t gj = fj > (Real)0 ? (fj < (Real)k.rows ? (uint32_t)fj : k.rows - 1) : 0u;
        uint32_t lx = gi - sv.col_off;
        if (lx >= sv.ncl) {                                  // a record outside the local grid: corrupt state, stop before indexing with it
            if (!atomicOr(&ds->flags, FLAG_CORRUPT)) { ds->dbg[0] = a; ds->dbg[1] = gi; ds->dbg[2] = gj; ds->dbg[3] = n; }
            lx = 0; gj = 0;
        }
        own.a_own = lx >= sv.gl && lx < sv.ncl - sv.gr;
        own.l_own = lx >= 1u && (lx - 1u) >= sv.gl && (lx - 1u) < sv.ncl - sv.gr;
        own.r_own = (lx + 1u) >= sv.gl && (lx + 1u) < sv.ncl - sv.gr;
        return load_stencil<Real>(cell_start, lx, gj, k.cols, k);
    }
    uint32_t gx, gy;
    cell_xy<Real>(pa.x, pa.y, k, gx, gy);
    return load_stencil<Real>(cell_start, gx, gy, k.cols, k);
}

// which of walk_contacts' entries this rank lists (all of them on one GPU)
__device__ __forceinline__ bool keep_entry(bool slab, const OwnFlags& own, uint32_t o, uint32_t centre) {
    return !slab || (centre ? own.a_own : (o >= 2u && o <= 3u ? own.l_own : (o == 4u ? own.r_own : own.a_own)));
}

template <class Real, bool SLAB>
__global__ void __launch_bounds__(DETECT_THREADS, (sizeof(Real) == 4 ? PS_DETECT_MINB : 4) * 256 / DETECT_THREADS)
detect_kernel(const Rec<Real>* __restrict__ rec, const uint32_t* __restrict__ cell_start, uint32_t n, uint32_t span, StepConst<Real> k,
              uint32_t* __restrict__ chunks, uint32_t* __restrict__ pool, uint32_t pool_cap, unsigned long long* __restrict__ head,
              uint2* __restrict__ wl0, DevState* __restrict__ ds, SlabView sv) {
    __shared__ uint2 queue[DETECT_THREADS / 32][DETECT_SPAN_MAX];     // per warp: {slot (+ DETECT_SLOW), hit mask}
    const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    const uint32_t below = (1u << lane) - 1u;
    uint2* const q = queue[warp];
    const uint32_t first = (blockIdx.x * (DETECT_THREADS / 32) + warp) * span;
    uint32_t n_blocked = 0u;
    if (SLAB) { n = *sv.n_local; n_blocked = *sv.n_blocked; }   // (`n` is the record capacity on entry; the slot count lives on the device)
    if (first >= n) return;
    // Error state (raised by an earlier launch, or by another warp of this one): no lists; the resolve kernel will not run.
    if (__any_sync(0xffffffffu, ds->flags != 0u)) return;
    const uint32_t end = first + span < n ? first + span : n;

    // ---- phase 1: test every candidate once, queue the slots that hit something
    uint32_t qn = 0u;                                        // queue length (same value in every lane)
    auto pa_next = first + lane < end ? ld_pos(rec + first + lane) : decltype(ld_pos(rec)){};   // the next batch's position is requested one batch ahead (1.100 -> 1.091 ms per sub-step at 16 Mi)
    for (uint32_t b0 = first; b0 < end; b0 += 32u) {
        const uint32_t a = b0 + lane;
        uint32_t hits = 0u, tag = a;
        const auto pa = pa_next;
        if (a + 32u < end) pa_next = ld_pos(rec + a + 32u);
        if (a < end) {
            OwnFlags own{true, true, true};
            const Stencil st = locate<Real, SLAB>(pa, cell_start, k, sv, ds, a, n, own);
            const uint32_t nL = st.l3 - st.l0, nM = st.m3 - st.m0;     // (all zero when the column does not exist)
            const uint32_t total = nL + nM + (st.r3 - st.r0);
            // list length bound: every candidate once, own-cell partners twice
            if (total + (st.m2 - st.m1) <= 32u) {
                hits = test_flat<Real>(rec, a, pa, st, nL, nM, total, k.r2sq);
            } else {
                hits = 1u; tag = a | DETECT_SLOW;            // phase 2 walks the block itself
            }
            if (SLAB) head[a] = 0ull;                        // slab_right_check_kernel looks at every slot of the last two columns
        }
        const uint32_t push = __ballot_sync(0xffffffffu, hits != 0u);
        if (hits) q[qn + (uint32_t)__popc(push & below)] = make_uint2(tag, hits);
        qn += (uint32_t)__popc(push);
    }
    __syncwarp();

    // ---- phase 2: the queued slots, full warps
    uint32_t wn = 0u;                                        // particles kept for the first work list, compacted at the front of the queue
    uint32_t my_entries = 0u;                                // statistics: entries written by this lane
    for (uint32_t i0 = 0; i0 < qn; i0 += 32u) {
        const bool live = i0 + lane < qn;
        const uint2 it = live ? q[i0 + lane] : make_uint2(0u, 0u);
        __syncwarp();                                        // everything of this batch is read before the compaction below writes
        const uint32_t a = it.x & ~DETECT_SLOW;
        bool listed = false;
        uint32_t e0 = 0u;
        if (live) {
            const auto pa = ld_pos(rec + a);
            OwnFlags own{true, true, true};
            const Stencil st = locate<Real, SLAB>(pa, cell_start, k, sv, ds, a, n, own);
            const bool blocked = SLAB && a < n_blocked;
            uint32_t* chunk = chunks + (size_t)a * LIST_CHUNK;
            uint32_t cnt = 0u, run = 0u;
            bool first_lower = false, fits = true;
            if (!(it.x & DETECT_SLOW)) {
                const uint32_t nL = st.l3 - st.l0, nM = st.m3 - st.m0;
                const ListRuns lr = list_runs(st, a);
                const uint32_t m32 = list_mask(it.y, st, a, nL, nM, lr, own);
                cnt = (uint32_t)__popc(m32);
                const bool ovf = cnt > LIST_CHUNK + 1u;
                if (ovf) {                                   // rare: a run in the pool for entries 8.. (one atomic per such particle)
                    run = atomicAdd(&ds->pool_used, cnt - LIST_CHUNK);
                    fits = (uint64_t)run + (cnt - LIST_CHUNK) <= pool_cap;
                    if (fits) chunk[LIST_CHUNK - 1u] = run;
                }
                if (cnt && fits) {
                    // Round 0 runs a ready pair from its lower slot (resolve_item), so a particle whose first partner has a
                    // lower slot has nothing to do there: it stays off the first work list.  Runs 3, 4, 6, 7 hold the higher slots.
                    const uint32_t p = (uint32_t)__ffs((int)m32) - 1u;
                    first_lower = !((p >= lr.t3 && p < lr.t5) || p >= lr.t6);
                    e0 = emit_list(m32, st, a, lr, chunk, pool, ovf, run) | (ovf ? HEAD_OVF : 0u);
                }
            } else {
                walk_contacts<Real>(a, rec, st, k.r2sq, [&](uint32_t, uint32_t o, uint32_t centre) { cnt += keep_entry(SLAB, own, o, centre) ? 1u : 0u; });
                const bool ovf = cnt > LIST_CHUNK + 1u;
                fits = cnt <= HEAD_REM_MAX;
                if (!fits) atomicOr(&ds->flags, FLAG_LIST_TOO_LONG);
                if (ovf && fits) {
                    run = atomicAdd(&ds->pool_used, cnt - LIST_CHUNK);
                    fits = (uint64_t)run + (cnt - LIST_CHUNK) <= pool_cap;
                    if (fits) chunk[LIST_CHUNK - 1u] = run;
                }
                if (cnt && fits) {
                    uint32_t j = 0u;
                    walk_contacts<Real>(a, rec, st, k.r2sq, [&](uint32_t qq, uint32_t o, uint32_t centre) {
                        if (!keep_entry(SLAB, own, o, centre)) return;
                        const uint32_t e = make_entry(qq, o, centre) | (j + 1u == cnt ? ENT_LAST : 0u);
                        if (j == 0u) { e0 = e; first_lower = qq < a; } else store_entry(chunk, pool, ovf, run, j, e);
                        ++j;
                    });
                    if (ovf) e0 |= HEAD_OVF;
                }
            }
            if (!fits) {                                     // does not fit: flag; the resolve kernel will not run
                atomicOr(&ds->flags, FLAG_CAPACITY);
                head[a] = 0ull;
            } else if (cnt) {
                head[a] = head_word(0u, 0u, e0 | (blocked ? 0u : HEAD_LIVE));
                if (SLAB && blocked) sv.brem[a] = cnt;       // brem is zero everywhere else (cleared at creation and by the unblock kernel)
                listed = !blocked && !first_lower;
                my_entries += cnt;
            } else if (!SLAB) {
                head[a] = 0ull;                              // (hits only with particles whose pair another rank lists: SLAB wrote 0 above)
            }
        }
        const uint32_t lb = __ballot_sync(0xffffffffu, listed);
        if (listed) q[wn + (uint32_t)__popc(lb & below)] = make_uint2(a, e0);
        wn += (uint32_t)__popc(lb);
    }
    __syncwarp();
    // ---- the span's piece of the first work list, and the entry count for the statistics
    my_entries = __reduce_add_sync(0xffffffffu, my_entries);
    uint32_t wbase = 0u;
    if (lane == 0) {
        if (my_entries) atomicAdd(&ds->entries_used, my_entries);
        if (wn) wbase = atomicAdd(&ds->wl_count[0], wn);
    }
    wbase = __shfl_sync(0xffffffffu, wbase, 0);
    for (uint32_t i = lane; i < wn; i += 32u) wl0[wbase + i] = q[i];
}


// Pair capture (ps_capture_pairs): the global ordered list {particle_i, particle_j} (src/core.zig:203).
template <class Real>
__global__ void __launch_bounds__(256)
count_centre_kernel(const Rec<Real>* __restrict__ rec, const uint32_t* __restrict__ cell_start, uint32_t n, StepConst<Real> k,
                    uint32_t* __restrict__ ctr_count) {
    const uint32_t a = blockIdx.x * blockDim.x + threadIdx.x;
    if (a >= n) return;
    const auto pa = ld_pos(rec + a);
    const Stencil st = load_stencil<Real>(cell_start, cell_of<Real>(pa.x, pa.y, k), k);
    uint32_t c = 0;
    walk_contacts<Real>(a, rec, st, k.r2sq, [&](uint32_t, uint32_t, uint32_t centre) { c += centre; });
    ctr_count[a] = c;
}

template <class Real>
__global__ void __launch_bounds__(256)
fill_pairs_kernel(const Rec<Real>* __restrict__ rec, const uint32_t* __restrict__ cell_start, const uint32_t* __restrict__ pid,
                  uint32_t n, StepConst<Real> k, const uint32_t* __restrict__ ctr_off, uint64_t* __restrict__ pairs,
                  uint32_t pairs_cap, const DevState* __restrict__ ds) {
    const uint32_t a = blockIdx.x * blockDim.x + threadIdx.x;
    if (a >= n || ds->ctr_total > pairs_cap) return;
    const auto pa = ld_pos(rec + a);
    const Stencil st = load_stencil<Real>(cell_start, cell_of<Real>(pa.x, pa.y, k), k);
    uint64_t w = ctr_off[a];
    const uint64_t id_a = pid[a];
    walk_contacts<Real>(a, rec, st, k.r2sq, [&](uint32_t q, uint32_t, uint32_t centre) {
        if (centre) { pairs[2 * w] = pid[q]; pairs[2 * w + 1] = id_a; ++w; }   // {neighbour, centre}
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
    const Real kk = (k.k_coef * van) * (Real)0.5;                             // :368  `/ 2.0`: halving is exact in binary floating point, so the
                                                                              //       product with 0.5 is the same bits as the IEEE quotient
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

// -DPS_TAIL_PROFILE (developer builds only, tools/tail_profile.sh): thread 0 of the tail CTA accumulates the clock cycles
// between marks along one tail round and prints the sums when the kernel ends.
#ifdef PS_TAIL_PROFILE
#define PS_MARK(i) do { if (LOCAL && threadIdx.x == 0) { const long long t__ = clock64(); prof[i] += t__ - prof[15]; prof[15] = t__; ++prof[14 - ((i) == 6)]; } } while (0)
#define PS_PROF_PARAM , long long (&prof)[16]
#define PS_PROF_ARG , prof
__device__ __forceinline__ void ps_keep(uint32_t x) { asm volatile("" :: "r"(x)); }
__device__ __forceinline__ void ps_keep(float x) { asm volatile("" :: "f"(x)); }
__device__ __forceinline__ void ps_keep(double x) { asm volatile("" :: "d"(x)); }
#define PS_KEEP(x) ps_keep(x)
#define PS_PROF_DUMMY long long prof[16] = {}; (void)prof;
#else
#define PS_PROF_DUMMY
#define PS_MARK(i) do {} while (0)
#define PS_PROF_PARAM
#define PS_PROF_ARG
#define PS_KEEP(x) do {} while (0)
#endif

constexpr uint32_t RESOLVE_CONTINUE = 0xffffffffu;
constexpr int RESOLVE_THREADS = 1024;     // one CTA per SM in the grid phase: the grid barrier has 148 participants
constexpr int RESOLVE_TAIL_THREADS = 256; // the single-CTA tail keeps 8 warps (a CTA barrier among 32 warps is slower)
constexpr uint32_t RESOLVE_TAIL = 512;    // work lists this short are finished by one CTA out of shared memory (deep chains: hundreds of short rounds)
constexpr uint32_t RESOLVE_TAIL_CAP = 2 * RESOLVE_TAIL;   // a round appends at most two items per item

// One work item {a, ea}: a particle whose head moved in the previous round (or any particle with a list when
// r = 0) together with its head entry.  Runs the pair if it is ready and this thread is the one to run it; returns
// how many particles advanced (0 or up to 2) with their new work items in `next`.
// LOCAL = false: several CTAs work on the round, other SMs' writes are read from L2 (ld.global.cg).
// LOCAL = true: one CTA alone (tail rounds); plain loads may be served by this SM's L1, which the CTA barrier between
// rounds keeps coherent with the CTA's own writes.  Chains revisit the same few particles round after round, so the
// L1 hits cut the per-round latency.  (Nothing was loaded through L1 earlier in the kernel, so no stale lines exist.)
template <class Real, bool LOCAL>
__device__ __forceinline__ int resolve_item(Rec<Real>* rec, const uint32_t* __restrict__ chunks, const uint32_t* __restrict__ pool,
                                            unsigned long long* head, uint2 item, uint32_t r, const StepConst<Real>& k, uint2 (&next)[2] PS_PROF_PARAM) {
    const uint32_t a = item.x, ea = item.y;
    const uint32_t b = ea & ENT_SLOT_MASK;
    // Round 0: every particle with a visible list is on the work list, and a pair can only be ready if the two
    // heads name each other, in which case the lower slot's thread runs it.  So the higher slot need not look.
    if (r == 0u && b < a) return 0;
    // everything that depends only on the item is requested at once (the records speculatively): one round trip to
    // memory decides whether the pair runs and brings what it needs
    const unsigned long long ha = LOCAL ? head[a] : __ldcg(head + a);
    const unsigned long long hb = LOCAL ? head[b] : __ldcg(head + b);
    const Rec<Real> Ra = LOCAL ? ld_rec(rec + a) : ld_rec_cg(rec + a), Rb = LOCAL ? ld_rec(rec + b) : ld_rec_cg(rec + b);
    if (head_stamp(ha) == r + 1) return 0;           // a already ran a pair this round (executed by its partner)
    PS_MARK(2);
    if (head_stamp(hb) == r + 1) return 0;           // b moved this round; b is on the next list and will look at a
    const uint32_t wa = head_low(ha), wb = head_low(hb);
    if (!(wb & HEAD_LIVE)) return 0;                 // (a is live: it is on the list)
    // same pair seen from the other side: partner = a, opposite role
    if ((wb & ENT_SLOT_MASK) != a || !((wb ^ ea) & ENT_CENTRE)) return 0;
    // Both particles on this round's list (b's head last moved in round r-1, or never when r = 0): the
    // lower slot runs the pair.  Membership is read from the SAME 64-bit word as the head entry, so a
    // concurrent executor cannot make the two threads disagree.
    if (b < a && head_stamp(hb) == r) return 0;
    PS_MARK(3);
    // the two particles' next entries (requested now, needed only when the head words are written back)
    const uint32_t pos_a = head_pos(ha), pos_b = head_pos(hb);
    const bool more_a = !(wa & ENT_LAST), more_b = !(wb & ENT_LAST);
    const uint32_t na = more_a ? (load_entry(chunks, pool, a, pos_a + 1u, wa & HEAD_OVF) | HEAD_LIVE | (wa & HEAD_OVF)) : 0u;
    const uint32_t nb = more_b ? (load_entry(chunks, pool, b, pos_b + 1u, wb & HEAD_OVF) | HEAD_LIVE | (wb & HEAD_OVF)) : 0u;
    PS_KEEP(na); PS_KEEP(nb); PS_MARK(4);
    const bool a_centre = (ea >> ENT_ROLE_BIT) & 1u; // entry = {particle_i = neighbour, particle_j = centre}
    Rec<Real> I = a_centre ? Rb : Ra, J = a_centre ? Ra : Rb;
    PS_KEEP(I.x); PS_KEEP(J.x); PS_MARK(5);
    resolve_pair<Real>(I, J, k);
    PS_KEEP(I.u); PS_KEEP(J.u); PS_KEEP(I.x); PS_MARK(6);
    st_rec(rec + (a_centre ? b : a), I);
    st_rec(rec + (a_centre ? a : b), J);
    // A particle whose sub-list ends with this pair is never looked at again in this sub-step (every partner's entry for it
    // has been consumed together with its own), and the tie-break above rests on stamps written before this round: its head
    // word need not be written.  The long rounds are bound by these small scattered stores, four per pair (measured:
    // fewer or wider loads, more loads in flight change nothing; dropping this one takes 12 % off the kernel).  Slab
    // handles keep the store for the last two local columns: slab_right_check_kernel reads their LIVE bits.
    if (LOCAL || more_a || a >= k.keep_from) head[a] = head_word(r + 1, pos_a + 1u, na);
    if (LOCAL || more_b || b >= k.keep_from) head[b] = head_word(r + 1, pos_b + 1u, nb);
    PS_MARK(7);
    const uint2 ia = make_uint2(a, na), ib = make_uint2(b, nb);     // (constant indices keep `next` in registers)
    next[0] = more_a ? ia : ib;
    next[1] = ib;
    return (more_a ? 1 : 0) + (more_b ? 1 : 0);
}

// The same work item for the tail rounds (one CTA, lists of a few dozen items, hundreds of rounds in a row): there the cost
// of a round is the dependent instruction chain of ONE warp, and every early exit above is a point where the warp splits and
// runs its pieces one after the other.  This version has no exits: every lane loads, tests and computes; one predicate decides
// who stores.  (Loads of a lane that will not run are harmless: every address is inside the arrays.)
template <class Real>
__device__ __forceinline__ int resolve_item_tail(Rec<Real>* rec, const uint32_t* __restrict__ chunks, const uint32_t* __restrict__ pool,
                                                 unsigned long long* head, uint2 item, uint32_t r, const StepConst<Real>& k, uint2 (&next)[2]) {
    const uint32_t a = item.x, ea = item.y;
    const uint32_t b = ea & ENT_SLOT_MASK;
    const unsigned long long ha = head[a], hb = head[b];
    const Rec<Real> Ra = ld_rec(rec + a), Rb = ld_rec(rec + b);
    const uint32_t wa = head_low(ha), wb = head_low(hb);
    const uint32_t pos_a = head_pos(ha), pos_b = head_pos(hb);
    // the words that hold the next entries when the lists go on inside their chunks (entry pos + 1 sits at chunk word pos)
    const uint32_t ca = chunks[(size_t)a * LIST_CHUNK + (pos_a < LIST_CHUNK - 1u ? pos_a : LIST_CHUNK - 1u)];
    const uint32_t cb = chunks[(size_t)b * LIST_CHUNK + (pos_b < LIST_CHUNK - 1u ? pos_b : LIST_CHUNK - 1u)];
    const bool go = !(r == 0u && b < a) && head_stamp(ha) != r + 1 && head_stamp(hb) != r + 1 && (wb & HEAD_LIVE) != 0u &&
                    (wb & ENT_SLOT_MASK) == a && ((wb ^ ea) & ENT_CENTRE) != 0u && !(b < a && head_stamp(hb) == r);
    const bool more_a = go && !(wa & ENT_LAST), more_b = go && !(wb & ENT_LAST);
    uint32_t na = ca, nb = cb;
    // lists longer than nine entries continue in the pool (rare: take the general path)
    if (more_a && (wa & HEAD_OVF) && pos_a >= LIST_CHUNK - 1u) na = load_entry(chunks, pool, a, pos_a + 1u, 1u);
    if (more_b && (wb & HEAD_OVF) && pos_b >= LIST_CHUNK - 1u) nb = load_entry(chunks, pool, b, pos_b + 1u, 1u);
    na = more_a ? (na | HEAD_LIVE | (wa & HEAD_OVF)) : 0u;
    nb = more_b ? (nb | HEAD_LIVE | (wb & HEAD_OVF)) : 0u;
    const bool a_centre = (ea >> ENT_ROLE_BIT) & 1u;
    Rec<Real> I = a_centre ? Rb : Ra, J = a_centre ? Ra : Rb;
    resolve_pair<Real>(I, J, k);
    if (go) {
        st_rec(rec + (a_centre ? b : a), I);
        st_rec(rec + (a_centre ? a : b), J);
        head[a] = head_word(r + 1, pos_a + 1u, na);
        head[b] = head_word(r + 1, pos_b + 1u, nb);
    }
    const uint2 ia = make_uint2(a, na), ib = make_uint2(b, nb);
    next[0] = more_a ? ia : ib;
    next[1] = ib;
    return (more_a ? 1 : 0) + (more_b ? 1 : 0);
}

// One round over the work list `wl[0..cnt)` by one or many CTAs; appends the particles advanced in this round, with
// their new head entry, to `out` (buffered per CTA: one global atomic per batch of 256 items).
template <class Real>
__device__ __forceinline__ void resolve_round(Rec<Real>* rec, const uint32_t* __restrict__ chunks, const uint32_t* __restrict__ pool, unsigned long long* head,
                                              const uint2* wl, uint32_t cnt, uint2* out, uint32_t* out_count, uint32_t r,
                                              uint32_t first, uint32_t stride, uint32_t cta_threads, const StepConst<Real>& k,
                                              uint2* sh_buf, uint32_t* sh_cnt, uint32_t* sh_base) {
    for (uint32_t base = first; base < cnt; base += stride) {
        if (threadIdx.x == 0) *sh_cnt = 0;
        __syncthreads();
        const uint32_t idx = base + threadIdx.x;
        if (idx < cnt) {
            uint2 next[2];
            PS_PROF_DUMMY
            const int m = resolve_item<Real, false>(rec, chunks, pool, head, __ldcg(wl + idx), r, k, next PS_PROF_ARG);
            for (int i = 0; i < m; ++i) sh_buf[atomicAdd(sh_cnt, 1u)] = next[i];
        }
        __syncthreads();
        const uint32_t m = *sh_cnt;
        if (m) {                                             // one global atomic per CTA per batch
            if (threadIdx.x == 0) *sh_base = atomicAdd(out_count, m);
            __syncthreads();
            for (uint32_t i = threadIdx.x; i < m; i += cta_threads) out[*sh_base + i] = sh_buf[i];
        }
        __syncthreads();
    }
}

// The same round for the grid phase, without any CTA barrier: every warp walks its share of the work list on its own
// (the next item is requested one batch ahead), collects the items it produces in its own shared-memory buffer and
// moves them to `out` with one global atomic per flush.  Warps therefore never wait for each other inside a round, and
// the chain of dependent L2 round trips of one batch overlaps with those of the other warps' batches.
constexpr uint32_t RESOLVE_WBUF = 128;   // items per warp buffer; a batch appends at most 64
template <class Real>
__device__ __forceinline__ void resolve_round_warps(Rec<Real>* rec, const uint32_t* __restrict__ chunks, const uint32_t* __restrict__ pool, unsigned long long* head,
                                                    const uint2* wl, uint32_t cnt, uint2* out, uint32_t* out_count, uint32_t r,
                                                    uint32_t first, uint32_t stride, const StepConst<Real>& k, uint2* sh_wbuf) {
    const uint32_t lane = threadIdx.x & 31u;
    uint2* buf = sh_wbuf + (threadIdx.x >> 5) * RESOLVE_WBUF;
    uint32_t held = 0;                                       // items in this warp's buffer (same value in every lane)
    auto flush = [&]() {
        uint32_t base = 0;
        if (lane == 0) base = atomicAdd(out_count, held);
        base = __shfl_sync(0xffffffffu, base, 0);
        __syncwarp();
        for (uint32_t i = lane; i < held; i += 32u) out[base + i] = buf[i];
        __syncwarp();
        held = 0;
    };
    uint32_t idx = first + threadIdx.x;                      // `first` is a multiple of 32: a warp's lanes share the trip count
    uint2 item = idx < cnt ? __ldcg(wl + idx) : make_uint2(0u, 0u);
    for (; idx - lane < cnt; idx += stride) {
        const bool live = idx < cnt;
        const uint2 cur = item;
        const uint32_t ahead = idx + stride;
        if (ahead < cnt) item = __ldcg(wl + ahead);
        uint2 next[2];
        PS_PROF_DUMMY
        const int m = live ? resolve_item<Real, false>(rec, chunks, pool, head, cur, r, k, next PS_PROF_ARG) : 0;
        const uint32_t one = __ballot_sync(0xffffffffu, m >= 1), two = __ballot_sync(0xffffffffu, m >= 2);
        if (one) {
            const uint32_t below = (1u << lane) - 1u;
            const uint32_t at = held + (uint32_t)__popc(one & below) + (uint32_t)__popc(two & below);
            if (m >= 1) buf[at] = next[0];
            if (m >= 2) buf[at + 1] = next[1];
            held += (uint32_t)__popc(one) + (uint32_t)__popc(two);
            if (held > RESOLVE_WBUF - 64u) flush();
        }
    }
    if (held) flush();
}

template <class Real>
__global__ void __launch_bounds__(RESOLVE_THREADS)
resolve_kernel(Rec<Real>* rec, const uint32_t* __restrict__ chunks, const uint32_t* __restrict__ pool, unsigned long long* head,
               uint2* wl0, uint2* wl1, uint2* wl2, StepConst<Real> k, DevState* ds, uint32_t r_start, const uint32_t* keep_from) {
    cg::grid_group grid = cg::this_grid();
    k.keep_from = keep_from ? *keep_from : 0xffffffffu;
    __shared__ uint2 sh_buf[(RESOLVE_THREADS / 32) * RESOLVE_WBUF];   // grid phase: one append buffer per warp; tail: one for the CTA
    __shared__ uint32_t sh_cnt, sh_base;
    auto list = [&](uint32_t i) { return i == 0 ? wl0 : (i == 1 ? wl1 : wl2); };   // (an indexed array would live in local memory)
    const bool leader = blockIdx.x == 0 && threadIdx.x == 0;
    if (ds->flags) {                                         // uniform across the grid: set before this launch
        if (leader) ds->wl_count[0] = 0;
        return;
    }
    // r_start > 0 only for the second phase of a slab sub-step (ps_slab.cuh); RESOLVE_CONTINUE = "the round after the last one
    // of the previous launch", read on the device so that the host need not know it
    uint32_t r = r_start == RESOLVE_CONTINUE ? ds->rounds_last + 1u : r_start;
    const uint32_t r_begin = r;
    const bool trace = leader && ds->trace_on && r_start == 0;
    const unsigned long long t_begin = trace ? global_ns() : 0ull;
    // grid phase: long work lists, one grid barrier per round
    for (;; ++r) {
        const uint32_t cur = r % 3, nxt = (r + 1) % 3, nn = (r + 2) % 3;
        const uint32_t cnt = __ldcg(&ds->wl_count[cur]);
        if (trace && r < 64) { ds->trace_cnt[r] = cnt; ds->trace_ns[r] = (uint32_t)(global_ns() - t_begin); }
        if (cnt <= RESOLVE_TAIL) break;                      // same value in every thread: the whole grid leaves together
        if (leader) ds->wl_count[nn] = 0;                    // consumed in round r-1, refilled in round r+1
        resolve_round_warps<Real>(rec, chunks, pool, head, list(cur), cnt, list(nxt), &ds->wl_count[nxt], r,
                                  blockIdx.x * RESOLVE_THREADS, gridDim.x * RESOLVE_THREADS, k, sh_buf);
        grid.sync();
    }
    if (blockIdx.x != 0 || threadIdx.x >= RESOLVE_TAIL_THREADS) return;     // exited threads count as arrived at later CTA barriers
    // Tail phase: CTA 0 alone.  A settled pile has hundreds of short rounds whose cost is the chain "read the two
    // particles - resolve - write them back - tell the other threads", so a round here costs ONE CTA barrier and nothing
    // else that is shared: the work list lives in shared memory (three rotating buffers carved out of the grid phase's
    // append buffers, with their lengths beside them), items are appended with one shared-memory atomic per executed pair, and
    // the particles' records and states are served by this SM's L1 (plain loads; the barrier makes the CTA's own
    // writes visible).  A list that outgrows the buffers (rare) goes back to global memory for a round.
#ifdef PS_TAIL_PROFILE
    constexpr bool LOCAL = true;
    long long prof[16] = {};
    prof[15] = clock64();
#endif
    uint2* const tl_items = sh_buf;                          // [3][RESOLVE_TAIL_CAP]
    __shared__ uint32_t tl_n[3];
    bool in_smem = false, rot_valid = false;
    uint32_t cur = 0, nxt = 0, nn = 0;
    for (;;) {
        if (!in_smem) {
            rot_valid = false;
            const uint32_t cur = r % 3, nxt = (r + 1) % 3, nn = (r + 2) % 3;
            const uint32_t cnt = __ldcg(&ds->wl_count[cur]);
            if (trace && r < 64) { ds->trace_cnt[r] = cnt; ds->trace_ns[r] = (uint32_t)(global_ns() - t_begin); }
            if (cnt == 0 || r + 2 >= HEAD_STAMP_MAX) break;
            __syncthreads();                                 // everyone has read the count before anything moves on
            if (cnt <= RESOLVE_TAIL) {                       // short enough: pull the list into shared memory
                for (uint32_t i = threadIdx.x; i < cnt; i += RESOLVE_TAIL_THREADS) tl_items[cur * RESOLVE_TAIL_CAP + i] = __ldcg(list(cur) + i);
                if (threadIdx.x == 0) { tl_n[cur] = cnt; tl_n[nxt] = 0; }
                in_smem = true;
                __syncthreads();
                continue;
            }
            if (threadIdx.x == 0) ds->wl_count[nn] = 0;
            resolve_round<Real>(rec, chunks, pool, head, list(cur), cnt, list(nxt), &ds->wl_count[nxt], r,
                                0u, RESOLVE_TAIL_THREADS, RESOLVE_TAIL_THREADS, k, sh_buf, &sh_cnt, &sh_base);
            __threadfence();
            __syncthreads();
            ++r;
            continue;
        }
        if (!rot_valid) { cur = r % 3; nxt = (r + 1) % 3; nn = (r + 2) % 3; rot_valid = true; }
        const uint32_t cnt = tl_n[cur];
        if (trace && r < 64) { ds->trace_cnt[r] = cnt; ds->trace_ns[r] = (uint32_t)(global_ns() - t_begin); }
        if (cnt == 0 || r + 2 >= HEAD_STAMP_MAX) break;
        if (cnt > RESOLVE_TAIL) {                            // grew beyond what the next round's appends could hold: back to global memory
            for (uint32_t i = threadIdx.x; i < cnt; i += RESOLVE_TAIL_THREADS) list(cur)[i] = tl_items[cur * RESOLVE_TAIL_CAP + i];
            if (threadIdx.x == 0) { ds->wl_count[cur] = cnt; ds->wl_count[nxt] = 0; }
            in_smem = false;
            __threadfence();
            __syncthreads();
            continue;
        }
        PS_MARK(0);
        if (threadIdx.x == 0) tl_n[nn] = 0;                  // the input of round r-1; everybody read it before that round's barrier
        const uint2* in = tl_items + cur * RESOLVE_TAIL_CAP;
        uint2* out = tl_items + nxt * RESOLVE_TAIL_CAP;
        for (uint32_t idx = threadIdx.x; idx < cnt; idx += RESOLVE_TAIL_THREADS) {
            uint2 next[2];
            PS_MARK(1);
#ifdef PS_TAIL_BRANCHY
            const int m = resolve_item<Real, true>(rec, chunks, pool, head, in[idx], r, k, next PS_PROF_ARG);
#else
            const int m = resolve_item_tail<Real>(rec, chunks, pool, head, in[idx], r, k, next);
#endif
            if (m) {                                         // (a shared-memory atomic per executed pair: no warp-wide vote, lanes stay independent)
                const uint32_t at = atomicAdd(&tl_n[nxt], (uint32_t)m);
                out[at] = next[0];
                if (m == 2) out[at + 1] = next[1];
            }
        }
        PS_MARK(8);
        __syncthreads();                                     // the round's one barrier: global writes and appended items become visible
        PS_MARK(9);
        ++r;
        { const uint32_t t = cur; cur = nxt; nxt = nn; nn = t; }      // (r % 3 without the divisions: this loop is one long dependent chain)
    }
#ifdef PS_TAIL_PROFILE
    if (threadIdx.x == 0 && ds->trace_on && r_start == 0)
        printf("[tail profile] rounds %u marks(top,item,heads,decided,entries,operands,computed,stored,appended,barrier) %lld %lld %lld %lld %lld %lld %lld %lld %lld %lld  passes %lld executed %lld\n",
               r, prof[0], prof[1], prof[2], prof[3], prof[4], prof[5], prof[6], prof[7], prof[8], prof[9], prof[14], prof[13]);
#endif
    uint32_t left = 0;
    if (threadIdx.x == 0) left = in_smem ? tl_n[r % 3] : __ldcg(&ds->wl_count[r % 3]);
    if (threadIdx.x == 0) {
        if (left != 0) atomicOr(&ds->flags, FLAG_CAPACITY);          // ran out of round stamps (never in practice)
        ds->rounds_last = r;
        if (trace) ds->trace_rounds = r < 63 ? r + 1 : 64;
        ds->rounds_total += r - r_begin;
        if (r_start == 0) {
            ds->inc_total = ds->entries_used;
            ds->pairs_total += ds->entries_used / 2;
        }
        ds->wl_count[0] = ds->wl_count[1] = ds->wl_count[2] = 0;
    }
}


// ---- host side ---------------------------------------------------------------------------------
// slots per detect warp: 256 for large scenes (few atomics, long coalesced work-list runs), shorter when that would
// leave SMs without work
inline uint32_t detect_span(uint64_t n) {
    uint64_t batches = n / (32ull * 148 * 48);
    if (batches < 1) batches = 1;
    if (batches > DETECT_SPAN_MAX / 32) batches = DETECT_SPAN_MAX / 32;
    return (uint32_t)batches * 32u;
}
inline unsigned detect_grid(uint64_t n, uint32_t span) {
    const uint64_t warps = (n + span - 1) / span;
    return (unsigned)((warps + DETECT_THREADS / 32 - 1) / (DETECT_THREADS / 32));
}

template <class Real>
struct Collider {
    uint64_t n = 0;
    uint32_t rows = 0, cols = 0, ncell = 0;
    bool capture = false;
    bool rec_valid = false;    // rec[cur]/pid[cur] hold the current state
    bool soa_valid = true;     // the caller's SoA arrays hold the current state
    int cur = 0;
    int device = 0;
    int resolve_grid = 0;

    uint32_t *cell_count = nullptr, *cell_start = nullptr, *scan_sums = nullptr;
    uint32_t *slot_cell = nullptr, *arrival = nullptr, *tmp_id = nullptr;
    Rec<Real>* rec[2] = {nullptr, nullptr};
    uint32_t* pid[2] = {nullptr, nullptr};
    uint32_t *ctr_count = nullptr, *ctr_off = nullptr;
    uint32_t* chunks = nullptr;                              // LIST_CHUNK entry words per slot
    uint32_t* entries = nullptr; uint32_t entries_cap = 0;   // overflow pool of the contact lists
    unsigned long long* head = nullptr;
    uint2* wl[3] = {nullptr, nullptr, nullptr};
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
        if ((rc = alloc(&cell_count, cell_count_bytes(ncell) / 4, err))) return rc;
        if ((rc = alloc(&cell_start, (uint64_t)ncell + 3, err))) return rc;
        if ((rc = alloc(&scan_sums, scan_tiles(scan_len) + 1, err))) return rc;
        if ((rc = alloc(&slot_cell, n, err))) return rc;
        if ((rc = alloc(&arrival, n, err))) return rc;
        if ((rc = alloc(&tmp_id, n, err))) return rc;
        for (int i = 0; i < 2; ++i) {
            if ((rc = alloc(&rec[i], n, err))) return rc;
            if ((rc = alloc(&pid[i], n, err))) return rc;
        }
        if ((rc = alloc(&entries, (uint64_t)entries_cap, err))) return rc;
        if ((rc = alloc(&head, n, err))) return rc;
        if ((rc = alloc(&chunks, (uint64_t)LIST_CHUNK * n, err))) return rc;
        for (int i = 0; i < 3; ++i) if ((rc = alloc(&wl[i], n, err))) return rc;
        if ((rc = alloc(&ds, 1, err))) return rc;
        PS_CUDA_TRY(err, cudaMemsetAsync(ds, 0, sizeof(DevState), st));
        if (getenv("PS_RESOLVE_TRACE")) { const uint32_t one = 1; PS_CUDA_TRY(err, cudaMemcpyAsync(&ds->trace_on, &one, sizeof(one), cudaMemcpyHostToDevice, st)); }
        int per_sm = 0, sms = 0;
        PS_CUDA_TRY(err, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, resolve_kernel<Real>, RESOLVE_THREADS, 0));
        PS_CUDA_TRY(err, cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device));
        int coop = 0;
        PS_CUDA_TRY(err, cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, device));
        if (!coop || per_sm < 1) { err.status = PS_ERR_CUDA; err.text = "device cannot run cooperative launches"; return err.status; }
        resolve_grid = per_sm * sms;
        return PS_OK;
    }

    void destroy() {
        cudaFree(cell_count); cudaFree(cell_start); cudaFree(scan_sums); cudaFree(slot_cell); cudaFree(arrival);
        cudaFree(tmp_id); cudaFree(ctr_count); cudaFree(ctr_off); cudaFree(entries); cudaFree(head); cudaFree(chunks);
        for (auto& r : rec) cudaFree(r);
        for (auto& p : pid) cudaFree(p);
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

    unsigned grid_n() const { return (unsigned)((n + 255) / 256); }

    // The caller wrote new SoA arrays (upload): the records are stale.
    void invalidate_records() { rec_valid = false; soa_valid = true; }
    // ... and whatever stopped the previous scene (capacity, cell overflow) no longer applies
    cudaError_t clear_flags(cudaStream_t st) { return cudaMemsetAsync(&ds->flags, 0, sizeof(uint32_t), st); }

    // Bring the SoA arrays (original index order) up to date with the records.
    int materialize_soa(Real* px, Real* py, Real* vx, Real* vy, cudaStream_t st, Error& err, uint64_t* launches) {
        if (soa_valid || !rec_valid || n == 0) { soa_valid = true; return PS_OK; }
        rec_to_soa_kernel<Real><<<grid_n(), 256, 0, st>>>(pid[cur], rec[cur], (uint32_t)n, px, py, vx, vy);
        PS_CUDA_TRY(err, cudaGetLastError());
        ++*launches;
        soa_valid = true;
        return PS_OK;
    }

    // One iteration of the loop body at src/core.zig:247-267.
    int substep(Real* px, Real* py, Real* vx, Real* vy, const StepConst<Real>& k, cudaStream_t st, Error& err, uint64_t* launches) {
        if (n == 0) return PS_OK;
        if (capture) { const int rc = ensure_capture(err); if (rc) return rc; }
        const uint32_t n32 = (uint32_t)n;
        const unsigned gn = grid_n();
        if (!rec_valid) {                                    // first sub-step after an upload: records in index order
            soa_to_rec_kernel<Real><<<gn, 256, 0, st>>>(px, py, vx, vy, n32, rec[cur], pid[cur]);
            rec_valid = true;
            ++*launches;
        }
        soa_valid = false;
        const int nw = cur ^ 1;
        // shg.clear() (:262) — only the counters need it; every other array is fully rewritten
        PS_CUDA_TRY(err, cudaMemsetAsync(cell_count, 0, cell_count_bytes(ncell), st));
        bin_kernel<Real><<<(gn + SORT_ILP - 1) / SORT_ILP, 256, 0, st>>>(rec[cur], n32, k, cell_count, slot_cell, arrival, ds);
        PS_CUDA_TRY(err, exclusive_scan<uint8_t>(reinterpret_cast<const uint8_t*>(cell_count), cell_start, ncell, scan_sums, cell_start + ncell, &ds->max_cell, st));
        place_kernel<<<(gn + SORT_ILP - 1) / SORT_ILP, 256, 0, st>>>(slot_cell, arrival, pid[cur], cell_start, n32, tmp_id);
        move_kernel<Real><<<(gn + SORT_ILP - 1) / SORT_ILP, 256, 0, st>>>(rec[cur], pid[cur], slot_cell, cell_start, tmp_id, n32, k, rec[nw], pid[nw], ds);
        {
            const uint32_t span = detect_span(n);
            detect_kernel<Real, false><<<detect_grid(n, span), DETECT_THREADS, 0, st>>>(rec[nw], cell_start, n32, span, k, chunks, entries, entries_cap, head, wl[0], ds, SlabView{});
        }
        *launches += 7;
        if (capture) {
            count_centre_kernel<Real><<<gn, 256, 0, st>>>(rec[nw], cell_start, n32, k, ctr_count);
            PS_CUDA_TRY(err, exclusive_scan_u32(ctr_count, ctr_off, n, scan_sums, &ds->ctr_total, nullptr, st));
            fill_pairs_kernel<Real><<<gn, 256, 0, st>>>(rec[nw], cell_start, pid[nw], n32, k, ctr_off, pairs, pairs_cap, ds);
            *launches += 5;
        }
        {
            Rec<Real>* a_rec = rec[nw]; const uint32_t* a_chunks = chunks; const uint32_t* a_ent = entries;
            unsigned long long* a_head = head; uint2 *a0 = wl[0], *a1 = wl[1], *a2 = wl[2];
            StepConst<Real> a_k = k; DevState* a_ds = ds; uint32_t a_r0 = 0; const uint32_t* a_keep = nullptr;
            void* args[] = {&a_rec, &a_chunks, &a_ent, &a_head, &a0, &a1, &a2, &a_k, &a_ds, &a_r0, &a_keep};
            const uint64_t want = (n + RESOLVE_THREADS - 1) / RESOLVE_THREADS;
            const unsigned grid = (unsigned)(want < (uint64_t)resolve_grid ? want : (uint64_t)resolve_grid);
            PS_CUDA_TRY(err, cudaLaunchCooperativeKernel((const void*)resolve_kernel<Real>, dim3(grid), dim3(RESOLVE_THREADS), args, 0, st));
            ++*launches;
        }
        PS_CUDA_TRY(err, cudaGetLastError());
        cur = nw;
        return PS_OK;
    }

    int check_flags(cudaStream_t st, Error& err) {
        DevState h{};
        PS_CUDA_TRY(err, cudaMemcpyAsync(&h, ds, sizeof(DevState), cudaMemcpyDeviceToHost, st));
        PS_CUDA_TRY(err, cudaStreamSynchronize(st));
        if (h.flags & FLAG_LIST_TOO_LONG) {
            err.status = PS_ERR_CAPACITY;
            err.text = "one particle has more than 8191 contacts (the per-particle limit of the cursor word); the scene is far denser than the reference's u8 cell counters allow";
            return err.status;
        }
        if (h.flags & FLAG_CAPACITY) {
            err.status = PS_ERR_CAPACITY;
            err.text = "the overflow pool of the contact lists (" + std::to_string(entries_cap) + " entries, for particles with more than 9 contacts) is full; raise ps_config.contacts_per_particle";
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
        if (h.trace_on) {                                    // PS_RESOLVE_TRACE: rounds of the last sub-step to stderr
            std::string line = "[resolve trace] round:items@us";
            for (uint32_t i = 0; i < h.trace_rounds; ++i) line += " " + std::to_string(i) + ":" + std::to_string(h.trace_cnt[i]) + "@" + std::to_string(h.trace_ns[i] / 1000);
            fprintf(stderr, "%s\n", line.c_str());
        }
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
