// ps_slab.cuh — the collision step slab-decomposed across the GPUs of one box.
//
// One process per GPU.  Rank k owns the particles whose cell column lies in [x0, x1); cells are
// column-major (src/core.zig:120), so a slab is a contiguous range of cell indices and the
// reference's global pair list is slab-major: every pair belongs to the rank that owns its CENTRE
// particle, and all of rank k's pairs precede all of rank k+1's.  Dependencies therefore cross a
// slab boundary in one direction only, through the particles of the two columns next to it:
//
//   M0  (before detection, both ways)  migrants (particles whose new column left the slab) and the
//       halo (own particles now in the boundary column), post-integrate state, with particle index.
//       Each rank sorts own + ghost columns into its local grid; both ranks of a boundary hold the
//       same particles of the two boundary columns in the same (cell, index) order.
//   M1  (rightward)  after rank k has executed its pairs: the state of its last own column and of
//       its right ghost column (= rank k+1's left ghost column and first own column, same order, so
//       the message is one contiguous range of records).  Until M1 arrives rank k+1 keeps those
//       particles "blocked": pairs that involve them wait, everything else proceeds (phase A); then
//       it unblocks them and finishes (phase B).
//   After M1 the particles of rank k's last own column are simply KEPT by rank k+1 (it holds their final state
//   as its left ghost column): ownership is handed over, rank k+1 integrates them at the next sub-step and routes
//   them by their new column — those still in rank k's columns travel back inside the next M0 as ordinary
//   migrants.  So there is no third message: a rank's owned slots are [0, start of its last own column).
//
// The result is bit-identical to the single-GPU / sequential reference order (the protocol is
// modelled on the CPU in tests/_slab_model.py and checked against the oracle).  Transport is NCCL
// ncclSend / ncclRecv between x-neighbours inside one ncclGroup per exchange (NVLink 5 / NVSwitch);
// messages are a few hundred KB, so the cost is latency: one host synchronisation per sub-step.
#pragma once
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "ps_collide.cuh"
#include "ps_nccl.h"

namespace ps {

constexpr uint32_t SLAB_INVALID = 0xffffffffu;
constexpr uint32_t SLAB_ERR_MIGRATION = 1u;   // a particle crossed a whole neighbouring slab in one sub-step
constexpr uint32_t SLAB_ERR_HALO = 2u;        // halo / migration buffer too small
constexpr uint32_t SLAB_ERR_CAPACITY = 4u;    // more local particles than the handle was sized for
constexpr uint32_t SLAB_ERR_PROTOCOL = 8u;    // boundary hand-over sizes disagree between ranks

struct SlabGeom {
    uint32_t rows, gcols;      // global grid
    uint32_t x0, x1;           // own global columns [x0, x1)
    uint32_t gl, gr;           // ghost column on the left / right
    uint32_t col_off, ncl;     // local grid: global column of local column 0, number of local columns
    uint32_t xlo, xhi;         // a particle of this rank may land in [xlo, xhi) (the neighbours' slabs)
    uint32_t cap, cap_h;       // capacities: local particles, records per halo buffer
};

struct SlabDev {               // device-side counters, copied to the host once per sub-step
    uint32_t n_local;          // slots in use (own + ghost columns)
    uint32_t own_b, own_e;     // owned slot range
    uint32_t c_first1, c_first2, c_last2, c_last1;   // cell_start at local columns 1, 2, ncl-2, ncl-1
    uint32_t n_blocked;        // = c_first2 when a left neighbour exists, else 0
    uint32_t right_incomplete; // some particle of the last two local columns still has pending pairs
    uint32_t err;
    uint32_t n_own_prev;       // owned particles at the start of this sub-step
    uint32_t zero;             // always 0 (range start for the pack / unpack kernels)
    uint32_t one;              // always 1
    uint32_t own_b_prev;       // first owned slot of the layout this sub-step started from
    uint32_t pad[2];
};

// halo / migration message: header + records + particle indices
struct HaloHeader { uint32_t count, pad0, pad1, pad2; };
template <class Real> __host__ __device__ inline size_t halo_bytes(uint32_t cap_h) { return sizeof(HaloHeader) + (size_t)cap_h * (sizeof(Rec<Real>) + sizeof(uint32_t)); }
template <class Real> __device__ __forceinline__ Rec<Real>* halo_recs(void* buf) { return reinterpret_cast<Rec<Real>*>(static_cast<char*>(buf) + sizeof(HaloHeader)); }
template <class Real> __device__ __forceinline__ uint32_t* halo_pids(void* buf, uint32_t cap_h) {
    return reinterpret_cast<uint32_t*>(static_cast<char*>(buf) + sizeof(HaloHeader) + (size_t)cap_h * sizeof(Rec<Real>));
}

template <class Real>
__device__ __forceinline__ void global_cell(Real x, Real y, const StepConst<Real>& k, uint32_t gcols, uint32_t& gi, uint32_t& gj) {
    const Real fi = floor_r(cell_coord<Real>(x, k)), fj = floor_r(cell_coord<Real>(y, k));   // computeCoordinate, src/core.zig:103-115
    gi = fi > (Real)0 ? (fi < (Real)gcols ? (uint32_t)fi : gcols - 1) : 0u;
    gj = fj > (Real)0 ? (fj < (Real)k.rows ? (uint32_t)fj : k.rows - 1) : 0u;
}

// ---- initial distribution -----------------------------------------------------------------------
// Keep the particles of this rank's columns out of a chunk of the global arrays (original index = first + i).
template <class Real>
__global__ void __launch_bounds__(256)
slab_filter_kernel(const double2* __restrict__ pos, const double2* __restrict__ vel, uint64_t first, uint32_t count,
                   StepConst<Real> k, SlabGeom g, Rec<Real>* __restrict__ rec, uint32_t* __restrict__ pid, SlabDev* __restrict__ sd) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= count) return;
    const double2 p = pos[i], v = vel[i];
    const Rec<Real> r{(Real)p.x, (Real)p.y, (Real)v.x, (Real)v.y};
    uint32_t gi, gj;
    global_cell<Real>(r.x, r.y, k, g.gcols, gi, gj);
    if (gi < g.x0 || gi >= g.x1) return;
    const uint32_t s = atomicAdd(&sd->own_e, 1u);
    if (s >= g.cap) { atomicOr(&sd->err, SLAB_ERR_CAPACITY); return; }
    st_rec(rec + s, r);
    pid[s] = (uint32_t)(first + i);
}

__device__ __forceinline__ uint64_t splitmix64_dev(uint64_t z) {
    z += 0x9E3779B97F4A7C15ull;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}
__device__ __forceinline__ double uniform01_dev(uint64_t base, uint64_t index, uint64_t stream) {
    return (double)(splitmix64_dev((index * 4ull + stream) ^ base) >> 11) * (1.0 / 9007199254740992.0);
}

// Spawner.spawnRandom's distribution (src/core.zig:403-408) with the counter-based generator of scenes.py,
// evaluated for every global index; a rank keeps what falls into its columns.  f64 arithmetic, multiply and
// add rounded separately, then cast: bit-identical to the host generator.
template <class Real>
__global__ void __launch_bounds__(256)
slab_spawn_kernel(uint64_t n_total, uint64_t seed_base, double radius, double width, double height, double vmin, double vmax,
                  StepConst<Real> k, SlabGeom g, Rec<Real>* __restrict__ rec, uint32_t* __restrict__ pid, SlabDev* __restrict__ sd) {
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_total; i += stride) {
        const double x = radius + (width - radius) * uniform01_dev(seed_base, i, 0);
        const Real rx = (Real)x;
        const Real fi = floor_r(cell_coord<Real>(rx, k));
        const uint32_t gi = fi > (Real)0 ? (fi < (Real)g.gcols ? (uint32_t)fi : g.gcols - 1) : 0u;
        if (gi < g.x0 || gi >= g.x1) continue;
        const double y = radius + (height - radius) * uniform01_dev(seed_base, i, 1);
        const double u = vmin + (vmax - vmin) * uniform01_dev(seed_base, i, 2);
        const double w = vmin + (vmax - vmin) * uniform01_dev(seed_base, i, 3);
        const uint32_t s = atomicAdd(&sd->own_e, 1u);
        if (s >= g.cap) { atomicOr(&sd->err, SLAB_ERR_CAPACITY); continue; }
        st_rec(rec + s, Rec<Real>{rx, (Real)y, (Real)u, (Real)w});
        pid[s] = (uint32_t)i;
    }
}

// The same generator writing a whole scene into SoA arrays in index order (single-GPU handle, ps_spawn).
template <class Real>
__global__ void __launch_bounds__(256)
spawn_soa_kernel(uint64_t n, uint64_t seed_base, double radius, double width, double height, double vmin, double vmax,
                 Real* __restrict__ px, Real* __restrict__ py, Real* __restrict__ vx, Real* __restrict__ vy) {
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        px[i] = (Real)(radius + (width - radius) * uniform01_dev(seed_base, i, 0));
        py[i] = (Real)(radius + (height - radius) * uniform01_dev(seed_base, i, 1));
        vx[i] = (Real)(vmin + (vmax - vmin) * uniform01_dev(seed_base, i, 2));
        vy[i] = (Real)(vmin + (vmax - vmin) * uniform01_dev(seed_base, i, 3));
    }
}

// ---- per sub-step kernels -----------------------------------------------------------------------
template <class Real>
__global__ void __launch_bounds__(256)
slab_begin_kernel(SlabDev* __restrict__ sd, DevState* __restrict__ ds, void* send_l, void* send_r) {
    if (threadIdx.x == 0 && blockIdx.x == 0) {
        sd->n_own_prev = sd->own_e - sd->own_b;
        sd->own_b_prev = sd->own_b;
        sd->right_incomplete = 0;
        ds->entries_used = 0;
        ds->pool_used = 0;
        reinterpret_cast<HaloHeader*>(send_l)->count = 0;
        reinterpret_cast<HaloHeader*>(send_r)->count = 0;
    }
}

// integrate + bounce, local cell + count for particles that stay in the local grid, migrants + halo into the send buffers
template <class Real>
__global__ void __launch_bounds__(256)
slab_bin_kernel(const Rec<Real>* __restrict__ rec_old, const uint32_t* __restrict__ pid_old, uint32_t own_b_prev, uint32_t n_own_prev,
                StepConst<Real> k, SlabGeom g,
                uint32_t* __restrict__ cell_count, uint32_t* __restrict__ slot_cell, uint32_t* __restrict__ arrival,
                void* send_l, void* send_r, SlabDev* __restrict__ sd, DevState* __restrict__ ds) {
    // SORT_ILP inputs per thread, CTA-strided: their loads, then their counting atomics, are in flight together.
    // The owned range is known on the host, so no thread waits for a count from memory first.
    const uint32_t t0 = blockIdx.x * (blockDim.x * SORT_ILP) + threadIdx.x;
    Rec<Real> r[SORT_ILP];
    uint32_t c[SORT_ILP], a[SORT_ILP];
#pragma unroll
    for (int j = 0; j < SORT_ILP; ++j) { const uint32_t t = t0 + j * blockDim.x; if (t < n_own_prev) r[j] = ld_rec(rec_old + own_b_prev + t); }
    const bool frozen = ds->flags != 0u;                             // error state: the state stays frozen
#pragma unroll
    for (int j = 0; j < SORT_ILP; ++j) {
        const uint32_t t = t0 + j * blockDim.x;
        c[j] = SLAB_INVALID;
        if (t >= n_own_prev) continue;
        const uint32_t s = own_b_prev + t;
        if (!frozen) integrate_rec<Real>(r[j], k);
        uint32_t gi, gj;
        global_cell<Real>(r[j].x, r[j].y, k, g.gcols, gi, gj);
        if (gi < g.xlo || gi >= g.xhi) atomicOr(&sd->err, SLAB_ERR_MIGRATION);
        if (g.gl && gi <= g.x0) {                                    // migrant (gi < x0) or halo (gi == x0) for the left rank
            const uint32_t q = atomicAdd(&reinterpret_cast<HaloHeader*>(send_l)->count, 1u);
            if (q < g.cap_h) { st_rec(halo_recs<Real>(send_l) + q, r[j]); halo_pids<Real>(send_l, g.cap_h)[q] = pid_old[s]; }
            else atomicOr(&sd->err, SLAB_ERR_HALO);
        }
        if (g.gr && gi + 1 >= g.x1) {                                // migrant (gi >= x1) or halo (gi == x1-1) for the right rank
            const uint32_t q = atomicAdd(&reinterpret_cast<HaloHeader*>(send_r)->count, 1u);
            if (q < g.cap_h) { st_rec(halo_recs<Real>(send_r) + q, r[j]); halo_pids<Real>(send_r, g.cap_h)[q] = pid_old[s]; }
            else atomicOr(&sd->err, SLAB_ERR_HALO);
        }
        if (gi >= g.col_off && gi < g.col_off + g.ncl) c[j] = (gi - g.col_off) * g.rows + gj;   // stays in the local grid (own or ghost column)
    }
#pragma unroll
    for (int j = 0; j < SORT_ILP; ++j) if (c[j] != SLAB_INVALID) a[j] = cell_count_fetch_inc(cell_count, c[j]);
#pragma unroll
    for (int j = 0; j < SORT_ILP; ++j) {
        const uint32_t t = t0 + j * blockDim.x;
        if (t >= n_own_prev) continue;
        slot_cell[t] = c[j];
        if (c[j] != SLAB_INVALID) {
            arrival[t] = a[j];
            if (a[j] >= 255u) atomicOr(&ds->flags, FLAG_CELL_OVERFLOW);
        }
    }
}

// received records (already integrated by their sender): input indices cap + j (left) and cap + cap_h + j (right)
template <class Real>
__global__ void __launch_bounds__(256)
slab_bin_recv_kernel(void* recv_l, void* recv_r, StepConst<Real> k, SlabGeom g, uint32_t* __restrict__ cell_count,
                     uint32_t* __restrict__ slot_cell, uint32_t* __restrict__ arrival, SlabDev* __restrict__ sd, DevState* __restrict__ ds) {
    const uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= 2 * g.cap_h) return;
    const bool right = j >= g.cap_h;
    if ((right && !g.gr) || (!right && !g.gl)) { slot_cell[g.cap + j] = SLAB_INVALID; return; }
    void* buf = right ? recv_r : recv_l;
    const uint32_t jj = right ? j - g.cap_h : j;
    const uint32_t cnt = reinterpret_cast<const HaloHeader*>(buf)->count;
    if (jj == 0 && cnt > g.cap_h) atomicOr(&sd->err, SLAB_ERR_HALO);
    if (jj >= cnt || jj >= g.cap_h) { slot_cell[g.cap + j] = SLAB_INVALID; return; }
    const Rec<Real> r = ld_rec(halo_recs<Real>(buf) + jj);
    uint32_t gi, gj;
    global_cell<Real>(r.x, r.y, k, g.gcols, gi, gj);
    if (gi >= g.col_off && gi < g.col_off + g.ncl) {
        const uint32_t c = (gi - g.col_off) * g.rows + gj;
        const uint32_t a = cell_count_fetch_inc(cell_count, c);
        slot_cell[g.cap + j] = c; arrival[g.cap + j] = a;
        if (a >= 255u) atomicOr(&ds->flags, FLAG_CELL_OVERFLOW);
    } else {
        slot_cell[g.cap + j] = SLAB_INVALID;                         // the neighbour's migrant deeper into its own slab: not ours to hold
    }
}

__global__ void slab_counts_kernel(const uint32_t* __restrict__ cell_start, SlabGeom g, SlabDev* __restrict__ sd) {
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    const uint32_t rows = g.rows, ncl = g.ncl;
    const uint32_t n = cell_start[ncl * rows];
    sd->n_local = n;
    if (n > g.cap) atomicOr(&sd->err, SLAB_ERR_CAPACITY);
    // owned after this sub-step: everything except the last own column and the right ghost column, which the right
    // rank keeps (it finishes their pairs); the left ghost column, finished here, is ours
    sd->own_b = 0u;
    sd->own_e = g.gr ? cell_start[(ncl >= 2 ? ncl - 2 : 0) * rows] : n;
    sd->c_first1 = cell_start[ncl >= 1 ? rows : 0];
    sd->c_first2 = cell_start[(ncl >= 2 ? 2 : ncl) * rows];
    sd->c_last2 = cell_start[(ncl >= 2 ? ncl - 2 : 0) * rows];
    sd->c_last1 = cell_start[(ncl >= 1 ? ncl - 1 : 0) * rows];
    sd->n_blocked = g.gl ? sd->c_first2 : 0u;
}

template <class Real>
__device__ __forceinline__ uint32_t slab_input_pid(uint32_t t, const uint32_t* pid_old, uint32_t own_b_prev, const SlabGeom& g, void* recv_l, void* recv_r) {
    if (t < g.cap) return pid_old[own_b_prev + t];
    const uint32_t j = t - g.cap;
    return j < g.cap_h ? halo_pids<Real>(recv_l, g.cap_h)[j] : halo_pids<Real>(recv_r, g.cap_h)[j - g.cap_h];
}

// input index space of place / move: [0, n_own_prev) own particles, then the 2 cap_h received records (stored at cap + j)
__device__ __forceinline__ uint32_t slab_input_index(uint32_t t, uint32_t n_own_prev, const SlabGeom& g) {
    if (t < n_own_prev) return t;
    const uint32_t j = t - n_own_prev;
    return j < 2 * g.cap_h ? g.cap + j : SLAB_INVALID;
}

// SORT_ILP inputs per thread, CTA-strided, the loads of each step issued together (as place_kernel / move_kernel do)
template <class Real>
__global__ void __launch_bounds__(256)
slab_place_kernel(const uint32_t* __restrict__ slot_cell, const uint32_t* __restrict__ arrival, const uint32_t* __restrict__ pid_old,
                  uint32_t own_b_prev, uint32_t n_own_prev, const uint32_t* __restrict__ cell_start, SlabGeom g, void* recv_l, void* recv_r,
                  uint32_t* __restrict__ tmp_id) {
    const uint32_t t0 = blockIdx.x * (blockDim.x * SORT_ILP) + threadIdx.x;
    uint32_t t[SORT_ILP], c[SORT_ILP], a[SORT_ILP], id[SORT_ILP], b[SORT_ILP];
#pragma unroll
    for (int j = 0; j < SORT_ILP; ++j) {
        t[j] = slab_input_index(t0 + j * blockDim.x, n_own_prev, g);
        c[j] = SLAB_INVALID;
        if (t[j] != SLAB_INVALID) { c[j] = slot_cell[t[j]]; a[j] = arrival[t[j]]; id[j] = slab_input_pid<Real>(t[j], pid_old, own_b_prev, g, recv_l, recv_r); }
    }
#pragma unroll
    for (int j = 0; j < SORT_ILP; ++j) if (c[j] != SLAB_INVALID) b[j] = cell_start[c[j]];
#pragma unroll
    for (int j = 0; j < SORT_ILP; ++j) {
        if (c[j] == SLAB_INVALID) continue;
        const uint32_t dst = b[j] + a[j];
        if (dst < g.cap) tmp_id[dst] = id[j];
    }
}

template <class Real>
__global__ void __launch_bounds__(256)
slab_move_kernel(const Rec<Real>* __restrict__ rec_old, const uint32_t* __restrict__ pid_old, uint32_t own_b_prev, uint32_t n_own_prev,
                 const uint32_t* __restrict__ slot_cell, const uint32_t* __restrict__ cell_start, const uint32_t* __restrict__ tmp_id,
                 StepConst<Real> k, SlabGeom g, void* recv_l, void* recv_r, const DevState* __restrict__ ds,
                 Rec<Real>* __restrict__ rec_new, uint32_t* __restrict__ pid_new) {
    const uint32_t t0 = blockIdx.x * (blockDim.x * SORT_ILP) + threadIdx.x;
    const bool frozen = ds->flags != 0u;
    uint32_t t[SORT_ILP], c[SORT_ILP], id[SORT_ILP], b[SORT_ILP], e[SORT_ILP];
    Rec<Real> r[SORT_ILP];
#pragma unroll
    for (int j = 0; j < SORT_ILP; ++j) {
        t[j] = slab_input_index(t0 + j * blockDim.x, n_own_prev, g);
        c[j] = SLAB_INVALID;
        if (t[j] == SLAB_INVALID) continue;
        c[j] = slot_cell[t[j]];
        id[j] = slab_input_pid<Real>(t[j], pid_old, own_b_prev, g, recv_l, recv_r);
        if (t[j] < g.cap) {
            r[j] = ld_rec(rec_old + own_b_prev + t[j]);
        } else {
            const uint32_t h = t[j] - g.cap;
            r[j] = h < g.cap_h ? ld_rec(halo_recs<Real>(recv_l) + h) : ld_rec(halo_recs<Real>(recv_r) + (h - g.cap_h));
        }
    }
#pragma unroll
    for (int j = 0; j < SORT_ILP; ++j) if (c[j] != SLAB_INVALID) { b[j] = cell_start[c[j]]; e[j] = cell_start[c[j] + 1]; }
#pragma unroll
    for (int j = 0; j < SORT_ILP; ++j) {
        if (c[j] == SLAB_INVALID) continue;
        uint32_t rank = 0;
        for (uint32_t q = b[j]; q < e[j] && q < g.cap; ++q) rank += tmp_id[q] < id[j] ? 1u : 0u;   // ascending particle index inside the cell (:146-158)
        if (t[j] < g.cap && !frozen) integrate_rec<Real>(r[j], k);              // same arithmetic as slab_bin_kernel, bit for bit; received records arrive integrated
        if (b[j] + rank >= g.cap) continue;                                      // only in an error state (capacity flagged by slab_counts_kernel)
        st_rec(rec_new + b[j] + rank, r[j]);
        pid_new[b[j] + rank] = id[j];
    }
}

// after phase A: do the particles of the last two local columns still have pending pairs?
__global__ void __launch_bounds__(256)
slab_right_check_kernel(const unsigned long long* __restrict__ head, const uint32_t* __restrict__ brem, SlabDev* __restrict__ sd) {
    const uint32_t a = sd->c_last2 + blockIdx.x * blockDim.x + threadIdx.x;
    if (a >= sd->n_local) return;
    if ((head_low(head[a]) & HEAD_LIVE) || brem[a] != 0u) sd->right_incomplete = 1u;
}

// M1 has been received into the records of the first two local columns: make their sub-lists visible
// (r0 = RESOLVE_CONTINUE: the round after phase A's last one, read from the device)
__global__ void __launch_bounds__(256)
slab_unblock_kernel(unsigned long long* __restrict__ head, uint32_t* __restrict__ brem, uint2* wl0, uint2* wl1, uint2* wl2, DevState* __restrict__ ds,
                    uint32_t r0, const SlabDev* __restrict__ sd) {
    const uint32_t a = blockIdx.x * blockDim.x + threadIdx.x;
    if (a >= sd->n_blocked) return;
    if (!brem[a]) return;
    brem[a] = 0u;
    if (r0 == RESOLVE_CONTINUE) r0 = ds->rounds_last + 1u;
    const uint32_t li = r0 % 3u;
    uint2* wl = li == 0u ? wl0 : (li == 1u ? wl1 : wl2);
    const uint32_t e = head_low(head[a]) | HEAD_LIVE;            // entry 0 as detect left it, now visible
    head[a] = head_word(r0, 0u, e);                              // stamp r0 = "on the work list of round r0"
    wl[atomicAdd(&ds->wl_count[li], 1u)] = make_uint2(a, e);
}

// boundary hand-over (M1): a contiguous range of records with its length in front; fixed-size messages,
// so that ranks that disagree about a count report SLAB_ERR_PROTOCOL instead of hanging inside NCCL
template <class Real>
__global__ void __launch_bounds__(256)
slab_pack_kernel(const Rec<Real>* __restrict__ rec, const uint32_t* __restrict__ first_p, const uint32_t* __restrict__ last_p,
                 uint32_t cap_h, void* buf, SlabDev* __restrict__ sd) {
    const uint32_t first = *first_p, n = *last_p - first;
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        reinterpret_cast<HaloHeader*>(buf)->count = n;
        if (n > cap_h) atomicOr(&sd->err, SLAB_ERR_HALO);
    }
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n && i < cap_h) st_rec(halo_recs<Real>(buf) + i, ld_rec_cg(rec + first + i));
}

template <class Real>
__global__ void __launch_bounds__(256)
slab_unpack_kernel(void* buf, Rec<Real>* __restrict__ rec, const uint32_t* __restrict__ first_p, const uint32_t* __restrict__ last_p,
                   uint32_t cap_h, SlabDev* __restrict__ sd) {
    const uint32_t first = *first_p, n = *last_p - first;
    const uint32_t got = reinterpret_cast<const HaloHeader*>(buf)->count;
    if (got != n) { if (blockIdx.x == 0 && threadIdx.x == 0) atomicOr(&sd->err, SLAB_ERR_PROTOCOL); return; }
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n && i < cap_h) st_rec(rec + first + i, ld_rec(halo_recs<Real>(buf) + i));
}

// ---- one-sided transport between x-neighbours (processes of one box, peer memory over NVLink) ----------------------
// A message is pushed straight into the neighbour's receive buffer (mapped with cudaIpcOpenMemHandle) by plain stores,
// followed by a system-wide fence and a sequence number in the neighbour's flag word; the neighbour's stream holds a
// one-warp kernel that waits for that number.  An exchange costs a few microseconds instead of the tens an
// ncclSend / ncclRecv pair takes, and it sits on the neighbour-to-neighbour critical cycle of every sub-step twice.
// `gate` (optional): send only if (*gate != 0) == gate_want — the boundary hand-over is queued twice, after phase A and
// after phase B, and the device decides which of the two carries it (no host round trip).
template <class Real>
__global__ void __launch_bounds__(1024)
slab_push_kernel(const void* __restrict__ src0, void* dst0, const void* __restrict__ src1, void* dst1, uint32_t cap_h,
                 uint32_t* flag0, uint32_t* flag1, uint32_t seq, const uint32_t* gate, uint32_t gate_want) {
    const void* src = blockIdx.x ? src1 : src0;
    void* dst = blockIdx.x ? dst1 : dst0;
    uint32_t* flag = blockIdx.x ? flag1 : flag0;
    if (!dst) return;
    if (gate && ((*gate != 0u) ? 1u : 0u) != gate_want) return;
    const uint32_t count = min(reinterpret_cast<const HaloHeader*>(src)->count, cap_h);
    const uint4* s4 = reinterpret_cast<const uint4*>(src);
    uint4* d4 = reinterpret_cast<uint4*>(dst);
    const uint32_t rec_v = (uint32_t)((sizeof(HaloHeader) + (size_t)count * sizeof(Rec<Real>)) / 16);       // header + records, in 16-byte units
    for (uint32_t i = threadIdx.x; i < rec_v; i += blockDim.x) d4[i] = s4[i];
    const size_t pid_off = (sizeof(HaloHeader) + (size_t)cap_h * sizeof(Rec<Real>)) / 16;                   // particle indices (slab_export needs them for M0)
    const uint32_t pid_v = (count + 3u) / 4u;
    for (uint32_t i = threadIdx.x; i < pid_v; i += blockDim.x) d4[pid_off + i] = s4[pid_off + i];
    __threadfence_system();
    __syncthreads();
    if (threadIdx.x == 0) { *reinterpret_cast<volatile uint32_t*>(flag) = seq; __threadfence_system(); }
}

// waits until the flag words hold `seq` (the neighbours' pushes of this sub-step have landed); gives up after ~20 s so that a
// dead neighbour turns into an error status instead of a kernel that never ends
__global__ void slab_wait_kernel(const uint32_t* flag0, const uint32_t* flag1, uint32_t seq, SlabDev* __restrict__ sd) {
    const uint32_t* flag = threadIdx.x == 0 ? flag0 : (threadIdx.x == 1 ? flag1 : nullptr);
    if (!flag) return;
    const unsigned long long t0 = global_ns();
    while (*reinterpret_cast<const volatile uint32_t*>(flag) != seq) {
        __nanosleep(200);
        if (global_ns() - t0 > 20000000000ull) { atomicOr(&sd->err, SLAB_ERR_PROTOCOL); return; }
    }
    __threadfence_system();
}

template <class Real>
__global__ void __launch_bounds__(256)
slab_export_kernel(const Rec<Real>* __restrict__ rec, const uint32_t* __restrict__ pid, uint32_t first, uint32_t count,
                   double2* __restrict__ pos, double2* __restrict__ vel, uint64_t* __restrict__ ids) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= count) return;
    const Rec<Real> r = ld_rec(rec + first + i);
    pos[i] = make_double2((double)r.x, (double)r.y);
    vel[i] = make_double2((double)r.u, (double)r.w);
    ids[i] = pid[first + i];
}

// ---- host side ------------------------------------------------------------------------------------
template <class Real>
struct SlabCollider {
    ps_config cfg{};
    int rank = 0, nranks = 1, device = 0;
    uint64_t n_total = 0;
    SlabGeom g{};
    uint32_t ncell_local = 0;
    NcclComm comm = nullptr;
    bool t_acc = false;
    uint64_t substeps_done = 0, launches = 0;
    cudaStream_t own_stream = nullptr, stream = nullptr;
    cudaEvent_t ev_rb = nullptr;                     // the mid-step copies of the device counters have landed
    Error err;

    int cur = 0;
    Rec<Real>* rec[2] = {nullptr, nullptr};
    uint32_t* pid[2] = {nullptr, nullptr};
    uint32_t *cell_count = nullptr, *cell_start = nullptr, *scan_sums = nullptr;
    uint32_t *slot_cell = nullptr, *arrival = nullptr, *tmp_id = nullptr, *brem = nullptr;
    uint32_t* chunks = nullptr;
    uint32_t* entries = nullptr; uint32_t entries_cap = 0;
    unsigned long long* head = nullptr;
    uint2* wl[3] = {nullptr, nullptr, nullptr};
    void *send_l = nullptr, *send_r = nullptr, *recv_l = nullptr, *recv_r = nullptr;
    void *in_l = nullptr, *in_r = nullptr, *in_m1 = nullptr;      // where this sub-step's M0 (from the left / right) and M1 have arrived
    // one-sided transport (see slab_push_kernel): my receive block, and the neighbours' blocks mapped into this process
    char* xbuf = nullptr; char *x_left = nullptr, *x_right = nullptr;
    bool p2p_l = false, p2p_r = false;
    size_t x_slot = 0;                                            // bytes per message buffer inside a receive block
    static constexpr size_t kFlagBytes = 1024;                    // six flag words, 128 bytes apart
    enum { X_M0_FROM_LEFT = 0, X_M0_FROM_RIGHT = 2, X_M1 = 4 };   // buffer / flag index = kind + parity of the sub-step
    uint32_t* x_flag(char* base, int idx) const { return reinterpret_cast<uint32_t*>(base + 128 * idx); }
    void* x_msg(char* base, int idx) const { return base + kFlagBytes + x_slot * idx; }
    DevState* ds = nullptr;
    SlabDev* sd = nullptr;
    struct HostCopy { SlabDev sd; DevState ds; };
    HostCopy* hc = nullptr;          // pinned
    SlabDev last{};                  // host copy after the most recent sub-step / load
    int resolve_grid = 0;
    void* stage = nullptr; size_t stage_bytes = 0;

    int fail(int status, const std::string& text) { err.status = status; err.text = text; return status; }
    // PS_SLAB_DEBUG=1: synchronise after every stage kernel and name the one that faulted
    bool debug_sync = getenv("PS_SLAB_DEBUG") != nullptr;
    bool force_late = getenv("PS_SLAB_FORCE_LATE") != nullptr;
    uint64_t late_sends = 0;       // sub-steps in which M1 had to wait for phase B
    int dbg(const char* what) {
        if (!debug_sync) return PS_OK;
        const cudaError_t e = cudaStreamSynchronize(stream);
        if (e != cudaSuccess) return fail(PS_ERR_CUDA, std::string("rank ") + std::to_string(rank) + " after " + what + ": " + cudaGetErrorString(e));
        return PS_OK;
    }
    // PS_SLAB_TIMING=1: CUDA-event time of every stage of the sub-step, summed and printed when the handle is destroyed
    bool timing = getenv("PS_SLAB_TIMING") != nullptr;
    static constexpr int kTicks = 9;
    static constexpr uint64_t kTimingSkip = 8;
    cudaEvent_t ev[kTicks] = {};
    bool ev_set[kTicks] = {};
    double acc_ms[kTicks] = {};
    uint64_t timing_from = 0;      // sub-steps done when timing was switched on
    void tick(int i) {
        if (!timing) return;
        if (!ev[i]) cudaEventCreate(&ev[i]);
        cudaEventRecord(ev[i], stream);
        ev_set[i] = true;
    }
    void tock() {
        if (!timing) return;
        cudaStreamSynchronize(stream);
        if (substeps_done <= timing_from + kTimingSkip) { for (int i = 0; i < kTicks; ++i) { acc_ms[i] = 0; ev_set[i] = false; } return; }   // connection set-up, first sort
        int prev = -1;
        for (int i = 0; i < kTicks; ++i) {
            if (!ev_set[i]) continue;
            if (prev >= 0) { float ms = 0; cudaEventElapsedTime(&ms, ev[prev], ev[i]); acc_ms[i] += ms; }
            prev = i; ev_set[i] = false;
        }
    }
    // ps_slab_set_timing / ps_slab_get_timing: the same figures through the C ABI (stages 1..8, microseconds per sub-step)
    void set_timing(bool on) {
        cudaStreamSynchronize(stream);
        timing = on;
        timing_from = substeps_done;
        for (int i = 0; i < kTicks; ++i) { acc_ms[i] = 0; ev_set[i] = false; }
    }
    uint64_t timed_substeps() const { return substeps_done > timing_from + kTimingSkip ? substeps_done - timing_from - kTimingSkip : 0; }
    void report() {
        if (!timing || timed_substeps() == 0) return;
        const double nsub = (double)timed_substeps();
        static const char* names[kTicks] = {"", "bin", "M0", "sort+detect", "resolve A+check", "readback", "M1", "phase B", "finish"};
        std::string line = "[slab timing] rank " + std::to_string(rank) + " per sub-step (us):";
        for (int i = 1; i < kTicks; ++i) line += std::string(" ") + names[i] + "=" + std::to_string((int)(acc_ms[i] * 1e3 / nsub));
        fprintf(stderr, "%s\n", line.c_str());
    }
    int nccl_try(int rc, const char* what) {
        if (rc == kNcclSuccess) return PS_OK;
        const NcclApi& api = nccl_api();
        return fail(PS_ERR_NCCL, std::string(what) + ": " + (api.GetErrorString ? api.GetErrorString(rc) : "NCCL error"));
    }
    template <class T> int alloc(T** p, uint64_t count) {
        PS_CUDA_TRY(err, cudaMalloc((void**)p, (count ? count : 1) * sizeof(T)));
        return PS_OK;
    }

    int create(const ps_config& c, int rank_, int nranks_, const NcclUniqueId* id, uint32_t rows, uint32_t cols) {
        cfg = c; rank = rank_; nranks = nranks_; device = c.device; n_total = c.count;
        PS_CUDA_TRY(err, cudaStreamCreateWithFlags(&own_stream, cudaStreamNonBlocking));
        stream = own_stream;
        PS_CUDA_TRY(err, cudaEventCreateWithFlags(&ev_rb, cudaEventDisableTiming));
        g.rows = rows; g.gcols = cols;
        g.x0 = (uint32_t)(((uint64_t)rank * cols) / nranks);
        g.x1 = (uint32_t)(((uint64_t)(rank + 1) * cols) / nranks);
        if (g.x1 - g.x0 < 2 && nranks > 1) return fail(PS_ERR_BAD_ARG, "ps_slab_create: every rank needs at least two cell columns");
        g.gl = rank > 0 ? 1u : 0u; g.gr = rank < nranks - 1 ? 1u : 0u;
        g.col_off = g.x0 - g.gl; g.ncl = (g.x1 - g.x0) + g.gl + g.gr;
        g.xlo = rank > 0 ? (uint32_t)(((uint64_t)(rank - 1) * cols) / nranks) : 0u;
        g.xhi = rank < nranks - 1 ? (uint32_t)(((uint64_t)(rank + 2) * cols) / nranks) : cols;
        const uint64_t share = (n_total + nranks - 1) / nranks;
        const uint64_t pct = c.slab_capacity_percent ? c.slab_capacity_percent : 125;
        uint64_t cap = n_total <= (4u << 20) ? n_total : share * pct / 100 + 16ull * rows + 4096;
        if (cap > n_total) cap = n_total;
        if (cap < 16) cap = 16;
        g.cap = (uint32_t)cap;
        g.cap_h = 4 * rows + 2048;
        if (g.cap_h > g.cap) g.cap_h = g.cap;
        ncell_local = g.ncl * rows;
        const uint64_t per = c.contacts_per_particle ? c.contacts_per_particle : 12;
        uint64_t ecap = per * cap; if (ecap > 0xfffffff0ull) ecap = 0xfffffff0ull;
        entries_cap = (uint32_t)ecap;
        int rc;
        for (int i = 0; i < 2; ++i) { if ((rc = alloc(&rec[i], cap))) return rc; if ((rc = alloc(&pid[i], cap))) return rc; }
        if ((rc = alloc(&cell_count, cell_count_bytes(ncell_local) / 4))) return rc;
        if ((rc = alloc(&cell_start, (uint64_t)ncell_local + 3))) return rc;
        if ((rc = alloc(&scan_sums, scan_tiles(ncell_local) + 1))) return rc;
        const uint64_t nin = cap + 2ull * g.cap_h;
        if ((rc = alloc(&slot_cell, nin))) return rc;
        if ((rc = alloc(&arrival, nin))) return rc;
        if ((rc = alloc(&tmp_id, cap))) return rc;
        if ((rc = alloc(&brem, cap))) return rc;
        if ((rc = alloc(&entries, (uint64_t)entries_cap))) return rc;
        if ((rc = alloc(&head, cap))) return rc;
        if ((rc = alloc(&chunks, (uint64_t)LIST_CHUNK * cap))) return rc;
        PS_CUDA_TRY(err, cudaMemsetAsync(head, 0, cap * sizeof(unsigned long long), stream));
        for (int i = 0; i < 3; ++i) if ((rc = alloc(&wl[i], cap))) return rc;
        const size_t hb = round_up(halo_bytes<Real>(g.cap_h), 256);     // (the push kernel copies whole 16-byte units)
        PS_CUDA_TRY(err, cudaMalloc(&send_l, hb)); PS_CUDA_TRY(err, cudaMalloc(&send_r, hb));
        PS_CUDA_TRY(err, cudaMalloc(&recv_l, hb)); PS_CUDA_TRY(err, cudaMalloc(&recv_r, hb));
        PS_CUDA_TRY(err, cudaMemsetAsync(send_l, 0, hb, stream)); PS_CUDA_TRY(err, cudaMemsetAsync(send_r, 0, hb, stream));
        PS_CUDA_TRY(err, cudaMemsetAsync(recv_l, 0, hb, stream)); PS_CUDA_TRY(err, cudaMemsetAsync(recv_r, 0, hb, stream));
        if ((rc = alloc(&ds, 1))) return rc;
        if ((rc = alloc(&sd, 1))) return rc;
        PS_CUDA_TRY(err, cudaMemsetAsync(ds, 0, sizeof(DevState), stream));
        PS_CUDA_TRY(err, cudaMemsetAsync(sd, 0, sizeof(SlabDev), stream));
        PS_CUDA_TRY(err, cudaMemsetAsync(brem, 0, cap * sizeof(uint32_t), stream));
        PS_CUDA_TRY(err, cudaHostAlloc((void**)&hc, sizeof(HostCopy), cudaHostAllocDefault));
        int per_sm = 0, sms = 0;
        PS_CUDA_TRY(err, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, resolve_kernel<Real>, RESOLVE_THREADS, 0));
        PS_CUDA_TRY(err, cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device));
        resolve_grid = per_sm * sms;
        in_l = recv_l; in_r = recv_r; in_m1 = recv_l;
        if (nranks > 1 && id) {                // id == NULL: ranks emulated in one process (SlabGroupLocal), no communicator
            const NcclApi& api = nccl_api();
            if (!api.ok) return fail(PS_ERR_NCCL, api.error);
            if ((rc = nccl_try(api.CommInitRank(&comm, nranks, *id, rank), "ncclCommInitRank"))) return rc;
            if ((rc = setup_peer_transport(hb))) return rc;
        }
        PS_CUDA_TRY(err, cudaStreamSynchronize(stream));
        return PS_OK;
    }

    // One-sided transport between x-neighbours: every rank allocates one receive block (flags + six message buffers), hands its
    // IPC handle to both neighbours over the communicator it already has, and maps theirs.  A boundary uses the one-sided path
    // only if BOTH of its ranks managed to map each other (they tell each other); otherwise it stays on ncclSend / ncclRecv.
    // PS_SLAB_TRANSPORT=nccl keeps every boundary on NCCL (set it on all ranks).
    int setup_peer_transport(size_t hb) {
        const char* want = getenv("PS_SLAB_TRANSPORT");
        const bool try_p2p = !(want && std::string(want) == "nccl");
        const NcclApi& api = nccl_api();
        x_slot = hb;
        struct Hello { cudaIpcMemHandle_t handle; uint32_t ok; uint32_t pad[3]; };
        Hello mine{}; Hello* dev = nullptr;
        if (try_p2p && cudaMalloc((void**)&xbuf, kFlagBytes + 6 * x_slot) == cudaSuccess) {
            PS_CUDA_TRY(err, cudaMemsetAsync(xbuf, 0, kFlagBytes + 6 * x_slot, stream));
            mine.ok = cudaIpcGetMemHandle(&mine.handle, xbuf) == cudaSuccess ? 1u : 0u;
        }
        cudaGetLastError();
        // round 1: handles; round 2: "I mapped yours"
        PS_CUDA_TRY(err, cudaMalloc((void**)&dev, 3 * sizeof(Hello)));                    // [0] mine, [1] from left, [2] from right
        Hello got[3]{};
        for (int round = 0; round < 2; ++round) {
            PS_CUDA_TRY(err, cudaMemcpyAsync(dev, &mine, sizeof(Hello), cudaMemcpyHostToDevice, stream));
            int rc;
            if ((rc = nccl_try(api.GroupStart(), "ncclGroupStart"))) return rc;
            if (g.gl) { api.Send(dev, sizeof(Hello), kNcclUint8, rank - 1, comm, stream); api.Recv(dev + 1, sizeof(Hello), kNcclUint8, rank - 1, comm, stream); }
            if (g.gr) { api.Send(dev, sizeof(Hello), kNcclUint8, rank + 1, comm, stream); api.Recv(dev + 2, sizeof(Hello), kNcclUint8, rank + 1, comm, stream); }
            if ((rc = nccl_try(api.GroupEnd(), "ncclGroupEnd (transport set-up)"))) return rc;
            PS_CUDA_TRY(err, cudaMemcpyAsync(got, dev, 3 * sizeof(Hello), cudaMemcpyDeviceToHost, stream));
            PS_CUDA_TRY(err, cudaStreamSynchronize(stream));
            if (round == 0) {
                // map the neighbours' blocks; `mine.ok` now carries, per side, whether that worked (bit 0 left, bit 1 right)
                uint32_t mapped = 0;
                if (g.gl && mine.ok && got[1].ok && cudaIpcOpenMemHandle((void**)&x_left, got[1].handle, cudaIpcMemLazyEnablePeerAccess) == cudaSuccess) mapped |= 1u;
                if (g.gr && mine.ok && got[2].ok && cudaIpcOpenMemHandle((void**)&x_right, got[2].handle, cudaIpcMemLazyEnablePeerAccess) == cudaSuccess) mapped |= 2u;
                cudaGetLastError();
                mine.ok = mapped;
            } else {
                p2p_l = g.gl && (mine.ok & 1u) && (got[1].ok & 2u);     // I mapped the left rank's block and it mapped mine (its right side)
                p2p_r = g.gr && (mine.ok & 2u) && (got[2].ok & 1u);
            }
        }
        cudaFree(dev);
        if (getenv("PS_SLAB_VERBOSE")) fprintf(stderr, "[slab] rank %d transport: left %s, right %s\n", rank, g.gl ? (p2p_l ? "peer stores" : "nccl") : "-", g.gr ? (p2p_r ? "peer stores" : "nccl") : "-");
        return PS_OK;
    }

    void destroy() {
        if (own_stream) cudaStreamSynchronize(own_stream);
        if (x_left) cudaIpcCloseMemHandle(x_left);
        if (x_right) cudaIpcCloseMemHandle(x_right);
        cudaFree(xbuf);
        if (ev_rb) cudaEventDestroy(ev_rb);
        report();
        if (comm) nccl_api().CommDestroy(comm);
        for (auto& r : rec) cudaFree(r);
        for (auto& p : pid) cudaFree(p);
        for (auto& w : wl) cudaFree(w);
        cudaFree(cell_count); cudaFree(cell_start); cudaFree(scan_sums); cudaFree(slot_cell); cudaFree(arrival);
        cudaFree(tmp_id); cudaFree(brem); cudaFree(entries); cudaFree(head); cudaFree(chunks);
        cudaFree(send_l); cudaFree(send_r); cudaFree(recv_l); cudaFree(recv_r); cudaFree(ds); cudaFree(sd); cudaFree(stage);
        if (hc) cudaFreeHost(hc);
        if (own_stream) cudaStreamDestroy(own_stream);
    }

    StepConst<Real> consts() const {
        StepConst<Real> k = make_step_const<Real>(cfg, t_acc, g.rows, g.gcols);
        k.cols = g.ncl;                 // the local grid is what the stencil walks; the global clamp uses g.gcols
        return k;
    }

    int ensure_stage(size_t bytes) {
        if (bytes <= stage_bytes) return PS_OK;
        cudaFree(stage); stage = nullptr; stage_bytes = 0;
        PS_CUDA_TRY(err, cudaMalloc(&stage, bytes));
        stage_bytes = bytes;
        return PS_OK;
    }

    int finish_load() {
        PS_CUDA_TRY(err, cudaMemcpyAsync(&hc->sd, sd, sizeof(SlabDev), cudaMemcpyDeviceToHost, stream));
        PS_CUDA_TRY(err, cudaStreamSynchronize(stream));
        if (hc->sd.err & SLAB_ERR_CAPACITY) return fail(PS_ERR_CAPACITY, "this rank's slab holds more particles than it was sized for");
        // slots [0, own_e) are all owned and in arbitrary order; the first sub-step sorts them
        SlabDev h = hc->sd;
        h.own_b = 0; h.n_local = h.own_e; h.one = 1u;
        PS_CUDA_TRY(err, cudaMemcpyAsync(sd, &h, sizeof(SlabDev), cudaMemcpyHostToDevice, stream));
        PS_CUDA_TRY(err, cudaStreamSynchronize(stream));
        last = h;
        return PS_OK;
    }

    int begin_load() {
        PS_CUDA_TRY(err, cudaMemsetAsync(sd, 0, sizeof(SlabDev), stream));
        PS_CUDA_TRY(err, cudaMemsetAsync(ds, 0, sizeof(DevState), stream));
        cur = 0;
        return PS_OK;
    }

    // every rank passes the same global arrays (n_total interleaved f64 pairs each) and keeps its own columns
    int upload_global(const double* pos_xy, const double* vel_xy) {
        int rc = begin_load(); if (rc) return rc;
        const StepConst<Real> k = consts();
        const uint64_t chunk = 4ull << 20;
        rc = ensure_stage(2 * chunk * sizeof(double2)); if (rc) return rc;
        double2* dpos = static_cast<double2*>(stage); double2* dvel = dpos + chunk;
        for (uint64_t first = 0; first < n_total; first += chunk) {
            const uint32_t m = (uint32_t)(n_total - first < chunk ? n_total - first : chunk);
            PS_CUDA_TRY(err, cudaMemcpyAsync(dpos, pos_xy + 2 * first, m * sizeof(double2), cudaMemcpyHostToDevice, stream));
            PS_CUDA_TRY(err, cudaMemcpyAsync(dvel, vel_xy + 2 * first, m * sizeof(double2), cudaMemcpyHostToDevice, stream));
            slab_filter_kernel<Real><<<(m + 255) / 256, 256, 0, stream>>>(dpos, dvel, first, m, k, g, rec[cur], pid[cur], sd);
            ++launches;
            PS_CUDA_TRY(err, cudaStreamSynchronize(stream));    // pageable source: keep the staging buffer safe
        }
        return finish_load();
    }

    int spawn(uint64_t seed, double vmin, double vmax) {
        int rc = begin_load(); if (rc) return rc;
        const StepConst<Real> k = consts();
        // splitmix64(seed) on the host = scenes.uniform01's `base`
        uint64_t z = seed + 0x9E3779B97F4A7C15ull;
        z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull; z = (z ^ (z >> 27)) * 0x94D049BB133111EBull; z ^= z >> 31;
        slab_spawn_kernel<Real><<<148 * 16, 256, 0, stream>>>(n_total, z, cfg.radius, cfg.width, cfg.height, vmin, vmax, k, g, rec[cur], pid[cur], sd);
        ++launches;
        PS_CUDA_TRY(err, cudaGetLastError());
        return finish_load();
    }

    int launch_resolve(const StepConst<Real>& k, uint32_t r0) {
        Rec<Real>* a_rec = rec[cur ^ 1]; const uint32_t* a_chunks = chunks; const uint32_t* a_ent = entries; unsigned long long* a_head = head;
        uint2 *a0 = wl[0], *a1 = wl[1], *a2 = wl[2]; StepConst<Real> a_k = k; DevState* a_ds = ds; uint32_t a_r0 = r0;
        const uint32_t* a_keep = &sd->c_last2;        // finished sub-lists of the last two local columns keep their head words current
        void* args[] = {&a_rec, &a_chunks, &a_ent, &a_head, &a0, &a1, &a2, &a_k, &a_ds, &a_r0, &a_keep};
        const uint64_t want = ((uint64_t)g.cap + RESOLVE_THREADS - 1) / RESOLVE_THREADS;
        const unsigned grid = (unsigned)(want < (uint64_t)resolve_grid ? want : (uint64_t)resolve_grid);
        PS_CUDA_TRY(err, cudaLaunchCooperativeKernel((const void*)resolve_kernel<Real>, dim3(grid), dim3(RESOLVE_THREADS), args, 0, stream));
        ++launches;
        return PS_OK;
    }

    int read_back() {
        PS_CUDA_TRY(err, cudaMemcpyAsync(&hc->sd, sd, sizeof(SlabDev), cudaMemcpyDeviceToHost, stream));
        PS_CUDA_TRY(err, cudaMemcpyAsync(&hc->ds, ds, sizeof(DevState), cudaMemcpyDeviceToHost, stream));
        PS_CUDA_TRY(err, cudaStreamSynchronize(stream));
        return check_host_copy();
    }

    // every device-side flag of this rank, from the host copies `hc` just read
    int check_host_copy() {
        const uint32_t e = hc->sd.err;
        if (e & SLAB_ERR_PROTOCOL) return fail(PS_ERR_NCCL, "boundary hand-over sizes disagree between neighbouring ranks");
        if (e & SLAB_ERR_MIGRATION) return fail(PS_ERR_MIGRATION, "a particle crossed a whole neighbouring slab in one sub-step");
        if (e & SLAB_ERR_HALO) return fail(PS_ERR_CAPACITY, "halo / migration buffer too small for this scene");
        if (e & SLAB_ERR_CAPACITY) return fail(PS_ERR_CAPACITY, "this rank's slab holds more particles than it was sized for");
        if (hc->ds.flags & FLAG_CORRUPT) return fail(PS_ERR_CUDA, "rank " + std::to_string(rank) + ": slot " + std::to_string(hc->ds.dbg[0]) + " of " + std::to_string(hc->ds.dbg[3]) +
                                                     " holds a record in global cell (" + std::to_string(hc->ds.dbg[1]) + "," + std::to_string(hc->ds.dbg[2]) + "), outside local columns [" +
                                                     std::to_string(g.col_off) + "," + std::to_string(g.col_off + g.ncl) + "); sd.err=" + std::to_string(hc->sd.err) +
                                                     " n_own_prev=" + std::to_string(hc->sd.n_own_prev) + " own=[" + std::to_string(hc->sd.own_b) + "," + std::to_string(hc->sd.own_e) + ")");
        if (hc->ds.flags & FLAG_LIST_TOO_LONG) return fail(PS_ERR_CAPACITY, "one particle has more than 8191 contacts (the per-particle limit of the head word); the scene is far denser than the reference's u8 cell counters allow");
        if (hc->ds.flags & FLAG_CAPACITY) return fail(PS_ERR_CAPACITY, "the overflow pool of the contact lists is full; raise ps_config.contacts_per_particle");
        if (hc->ds.flags & FLAG_CELL_OVERFLOW) return fail(PS_ERR_CELL_OVERFLOW, "a grid cell held more than 255 particles (src/core.zig:51,132)");
        return PS_OK;
    }

    // ---- one sub-step (the loop body at src/core.zig:247-267, this rank's share) in stages ------------------
    // The stages are separate so that the same code runs with NCCL between processes (substep()) or with plain
    // device copies between several ranks emulated in one process on one GPU (SlabGroupLocal, used by the tests).
    StepConst<Real> k_step{};
    SlabDev h_step{};
    uint32_t r0_step = 0;
    bool send_now_step = false;
    uint32_t own_b_step = 0, n_own_step = 0;
    size_t halo_msg_bytes() const { return halo_bytes<Real>(g.cap_h); }

    // stage 1: integrate + bin own particles; send_l / send_r now hold the M0 messages
    int stage_bin() {
        if (err.status != PS_OK) return err.status;
        k_step = consts();
        cudaStream_t st = stream;
        PS_CUDA_TRY(err, cudaMemsetAsync(cell_count, 0, cell_count_bytes(ncell_local), st));
        slab_begin_kernel<Real><<<1, 32, 0, st>>>(sd, ds, send_l, send_r);
        own_b_step = last.own_b; n_own_step = last.own_e - last.own_b;   // the layout left by the previous sub-step (host copy)
        slab_bin_kernel<Real><<<(n_own_step + 256 * SORT_ILP - 1) / (256 * SORT_ILP) + 1, 256, 0, st>>>(rec[cur], pid[cur], own_b_step, n_own_step, k_step, g, cell_count, slot_cell, arrival, send_l, send_r, sd, ds);
        launches += 2;
        PS_CUDA_TRY(err, cudaGetLastError());
        return PS_OK;
    }

    // stage 2 (recv_l / recv_r hold the neighbours' M0): local sort, detect, phase A, one host synchronisation
    int stage_build() {
        const int rc = stage_build_queue();
        return rc ? rc : stage_build_wait();
    }
    int stage_build_queue() {
        const int nw = cur ^ 1;
        const unsigned g_in = (n_own_step + 2 * g.cap_h + 256 * SORT_ILP - 1) / (256 * SORT_ILP), g_h = (2 * g.cap_h + 255) / 256;
        cudaStream_t st = stream;
        const StepConst<Real>& k = k_step;
        if (int d = dbg("bin")) return d;
        slab_bin_recv_kernel<Real><<<g_h, 256, 0, st>>>(in_l, in_r, k, g, cell_count, slot_cell, arrival, sd, ds);
        if (int d = dbg("bin_recv")) return d;
        PS_CUDA_TRY(err, exclusive_scan<uint8_t>(reinterpret_cast<const uint8_t*>(cell_count), cell_start, ncell_local, scan_sums, cell_start + ncell_local, &ds->max_cell, st));
        if (int d = dbg("scan")) return d;
        slab_counts_kernel<<<1, 32, 0, st>>>(cell_start, g, sd);
        if (int d = dbg("counts")) return d;
        slab_place_kernel<Real><<<g_in, 256, 0, st>>>(slot_cell, arrival, pid[cur], own_b_step, n_own_step, cell_start, g, in_l, in_r, tmp_id);
        if (int d = dbg("place")) return d;
        slab_move_kernel<Real><<<g_in, 256, 0, st>>>(rec[cur], pid[cur], own_b_step, n_own_step, slot_cell, cell_start, tmp_id, k, g, in_l, in_r, ds, rec[nw], pid[nw]);
        if (int d = dbg("move")) return d;
        SlabView sv{g.gl, g.gr, g.ncl, g.col_off, g.gcols, &sd->n_local, &sd->n_blocked, brem};
        const uint64_t n_hi = (uint64_t)n_own_step + 2ull * g.cap_h;             // at most this many local slots after the exchange
        const uint64_t n_det = n_hi < g.cap ? n_hi : g.cap;
        const uint32_t span = detect_span(n_det);
        detect_kernel<Real, true><<<detect_grid(n_det, span), DETECT_THREADS, 0, st>>>(rec[nw], cell_start, g.cap, span, k, chunks, entries, entries_cap, head, wl[0], ds, sv);
        launches += 9;
        int rc;
        if (int d = dbg("detect")) return d;
        tick(3);
        if ((rc = launch_resolve(k, 0u))) return rc;                              // phase A
        if (int d = dbg("resolve A")) return d;
        if (g.gr) { slab_right_check_kernel<<<(2 * g.cap_h + 255) / 256, 256, 0, st>>>(head, brem, sd); ++launches; }   // two columns of slots at most
        tick(4);
        PS_CUDA_TRY(err, cudaMemcpyAsync(&hc->sd, sd, sizeof(SlabDev), cudaMemcpyDeviceToHost, stream));
        PS_CUDA_TRY(err, cudaMemcpyAsync(&hc->ds, ds, sizeof(DevState), cudaMemcpyDeviceToHost, stream));
        PS_CUDA_TRY(err, cudaEventRecord(ev_rb, stream));
        return PS_OK;
    }
    // ... and the sub-step's one host synchronisation: on those copies only, not on what was queued after them
    int stage_build_wait() {
        int rc;
        PS_CUDA_TRY(err, cudaEventSynchronize(ev_rb));
        if ((rc = check_host_copy())) return rc;
        h_step = hc->sd;
        r0_step = hc->ds.rounds_last + 1;
        // PS_SLAB_FORCE_LATE=1 (tests): behave as if the right boundary always depended on the left one
        send_now_step = g.gr && !h_step.right_incomplete && !force_late;
        if (g.gr && !send_now_step) ++late_sends;
        return PS_OK;
    }

    // M1 out: last own column + right ghost column -> send_r (they are the right rank's left ghost + first own column)
    int pack_m1() {
        slab_pack_kernel<Real><<<(g.cap_h + 255) / 256, 256, 0, stream>>>(rec[cur ^ 1], &sd->c_last2, &sd->n_local, g.cap_h, send_r, sd);
        ++launches;
        PS_CUDA_TRY(err, cudaGetLastError());
        return PS_OK;
    }

    // stage 3 (recv_l holds the left rank's M1): take over the boundary states, unblock, phase B
    int stage_phase_b(uint32_t r0) {
        cudaStream_t st = stream;
        slab_unpack_kernel<Real><<<(g.cap_h + 255) / 256, 256, 0, st>>>(in_m1, rec[cur ^ 1], &sd->zero, &sd->c_first2, g.cap_h, sd);
        slab_unblock_kernel<<<(2 * g.cap_h + 255) / 256 + 1, 256, 0, st>>>(head, brem, wl[0], wl[1], wl[2], ds, r0, sd);   // two columns of slots at most
        launches += 2;
        if (int d = dbg("unpack+unblock")) return d;
        if (int rc = launch_resolve(k_step, r0)) return rc;
        return dbg("resolve B");
    }
    int stage_phase_b() { return stage_phase_b(r0_step); }

    // stage 4: finish the sub-step (the new layout becomes current)
    int stage_finish() {
        PS_CUDA_TRY(err, cudaGetLastError());
        cur ^= 1;
        last = h_step;
        ++substeps_done;
        return PS_OK;
    }

    int push(const void* src0, void* dst0, uint32_t* flag0, const void* src1, void* dst1, uint32_t* flag1, uint32_t seq, const uint32_t* gate, uint32_t gate_want) {
        slab_push_kernel<Real><<<2, 1024, 0, stream>>>(src0, dst0, src1, dst1, g.cap_h, flag0, flag1, seq, gate, gate_want);
        ++launches;
        PS_CUDA_TRY(err, cudaGetLastError());
        return PS_OK;
    }

    // One process per GPU: the stages above with the exchanges between x-neighbours — one-sided stores into the neighbour's
    // receive block where both sides could map each other (the usual case on one NVSwitch box), ncclSend / ncclRecv otherwise.
    int substep() {
        const NcclApi& api = nccl_api();
        const size_t hb = halo_msg_bytes();
        cudaStream_t st = stream;
        const uint32_t seq = (uint32_t)(substeps_done + 1);
        const int par = (int)(seq & 1u);
        int rc;
        tick(0);
        if ((rc = stage_bin())) return rc;
        tick(1);
        if (nranks > 1) {                                                         // ---- M0: migrants + halo, both ways
            if (p2p_l || p2p_r) {
                if ((rc = push(send_l, p2p_l ? x_msg(x_left, X_M0_FROM_RIGHT + par) : nullptr, p2p_l ? x_flag(x_left, X_M0_FROM_RIGHT + par) : nullptr,
                               send_r, p2p_r ? x_msg(x_right, X_M0_FROM_LEFT + par) : nullptr, p2p_r ? x_flag(x_right, X_M0_FROM_LEFT + par) : nullptr,
                               seq, nullptr, 0u))) return rc;
            }
            if ((g.gl && !p2p_l) || (g.gr && !p2p_r)) {
                if ((rc = nccl_try(api.GroupStart(), "ncclGroupStart"))) return rc;
                if (g.gl && !p2p_l) { api.Send(send_l, hb, kNcclUint8, rank - 1, comm, st); api.Recv(recv_l, hb, kNcclUint8, rank - 1, comm, st); }
                if (g.gr && !p2p_r) { api.Send(send_r, hb, kNcclUint8, rank + 1, comm, st); api.Recv(recv_r, hb, kNcclUint8, rank + 1, comm, st); }
                if ((rc = nccl_try(api.GroupEnd(), "ncclGroupEnd (M0)"))) return rc;
            }
            if (p2p_l || p2p_r) {
                slab_wait_kernel<<<1, 32, 0, st>>>(p2p_l ? x_flag(xbuf, X_M0_FROM_LEFT + par) : nullptr, p2p_r ? x_flag(xbuf, X_M0_FROM_RIGHT + par) : nullptr, seq, sd);
                ++launches;
            }
            in_l = p2p_l ? x_msg(xbuf, X_M0_FROM_LEFT + par) : recv_l;
            in_r = p2p_r ? x_msg(xbuf, X_M0_FROM_RIGHT + par) : recv_r;
        }
        tick(2);
        if ((rc = stage_build_queue())) return rc;
        // With one-sided transport on every boundary of this rank nothing that follows needs a decision of the host (the device
        // picks the early or late hand-over, phase B continues from the round phase A ended with): the rest of the sub-step is
        // queued first and the host looks at the mid-step counters afterwards, while the GPU is busy with the hand-over.
        const bool queue_ahead = (!g.gl || p2p_l) && (!g.gr || p2p_r) && !debug_sync;
        if (!queue_ahead && (rc = stage_build_wait())) return rc;
        tick(5);
        const bool send_now = send_now_step;
        if (g.gr && p2p_r) {
            // hand-over to the right, queued twice (here and after phase B); the device sends the one that applies
            if ((rc = pack_m1())) return rc;
            if ((rc = push(send_r, x_msg(x_right, X_M1 + par), x_flag(x_right, X_M1 + par), nullptr, nullptr, nullptr, seq, force_late ? &sd->one : &sd->right_incomplete, 0u))) return rc;
        } else if (send_now && (rc = pack_m1())) return rc;
        if (nranks > 1 && ((g.gr && !p2p_r && send_now) || (g.gl && !p2p_l))) {     // ---- M1: boundary states to the right
            if ((rc = nccl_try(api.GroupStart(), "ncclGroupStart"))) return rc;
            if (g.gr && !p2p_r && send_now) api.Send(send_r, hb, kNcclUint8, rank + 1, comm, st);
            if (g.gl && !p2p_l) api.Recv(recv_l, hb, kNcclUint8, rank - 1, comm, st);
            if ((rc = nccl_try(api.GroupEnd(), "ncclGroupEnd (M1)"))) return rc;
        }
        if (g.gl && p2p_l) {
            slab_wait_kernel<<<1, 32, 0, st>>>(x_flag(xbuf, X_M1 + par), nullptr, seq, sd);
            ++launches;
        }
        in_m1 = p2p_l ? x_msg(xbuf, X_M1 + par) : recv_l;
        tick(6);
        if (g.gl && (rc = stage_phase_b(queue_ahead ? RESOLVE_CONTINUE : r0_step))) return rc;
        tick(7);
        if (g.gr && p2p_r) {
            if ((rc = pack_m1())) return rc;
            if ((rc = push(send_r, x_msg(x_right, X_M1 + par), x_flag(x_right, X_M1 + par), nullptr, nullptr, nullptr, seq, force_late ? &sd->one : &sd->right_incomplete, 1u))) return rc;
        } else if (g.gr && !send_now) {                                           // rare: our right boundary depended on the left one
            if ((rc = pack_m1())) return rc;
            if ((rc = nccl_try(api.Send(send_r, hb, kNcclUint8, rank + 1, comm, st), "ncclSend (late M1)"))) return rc;
        }
        if (queue_ahead && (rc = stage_build_wait())) return rc;
        rc = stage_finish();
        tick(8);
        tock();
        return rc;
    }

    // Waits for everything queued and reports what the LAST sub-step raised after its one mid-step read-back (phase B's
    // resolve, a late hand-over): without this a failed final sub-step would look like success.
    int sync() {
        if (err.status != PS_OK) return err.status;
        return read_back();
    }

    // this rank's owned particles: global indices + interleaved f64 position / velocity
    int download_owned(uint64_t* ids, double* pos_xy, double* vel_xy, uint64_t cap_out, uint64_t* count) {
        PS_CUDA_TRY(err, cudaStreamSynchronize(stream));
        const uint32_t n = last.own_e - last.own_b;
        *count = n;
        if (n == 0) return PS_OK;
        if (cap_out < n) return fail(PS_ERR_BAD_ARG, "ps_slab_download_owned: output arrays too small");
        int rc = ensure_stage((size_t)n * (2 * sizeof(double2) + sizeof(uint64_t))); if (rc) return rc;
        double2* dpos = static_cast<double2*>(stage); double2* dvel = dpos + n; uint64_t* dids = reinterpret_cast<uint64_t*>(dvel + n);
        slab_export_kernel<Real><<<(n + 255) / 256, 256, 0, stream>>>(rec[cur], pid[cur], last.own_b, n, dpos, dvel, dids);
        ++launches;
        if (pos_xy) PS_CUDA_TRY(err, cudaMemcpyAsync(pos_xy, dpos, (size_t)n * sizeof(double2), cudaMemcpyDeviceToHost, stream));
        if (vel_xy) PS_CUDA_TRY(err, cudaMemcpyAsync(vel_xy, dvel, (size_t)n * sizeof(double2), cudaMemcpyDeviceToHost, stream));
        if (ids) PS_CUDA_TRY(err, cudaMemcpyAsync(ids, dids, (size_t)n * sizeof(uint64_t), cudaMemcpyDeviceToHost, stream));
        PS_CUDA_TRY(err, cudaStreamSynchronize(stream));
        return PS_OK;
    }

    int stats(ps_stats* out) {
        PS_CUDA_TRY(err, cudaMemcpyAsync(&hc->ds, ds, sizeof(DevState), cudaMemcpyDeviceToHost, stream));
        PS_CUDA_TRY(err, cudaStreamSynchronize(stream));
        std::memset(out, 0, sizeof(*out));
        out->substeps_done = substeps_done;
        out->pairs_last = hc->ds.inc_total / 2;
        out->pairs_total = hc->ds.pairs_total;
        out->rounds_last = hc->ds.rounds_last;
        out->rounds_total = hc->ds.rounds_total;
        out->max_cell_count = hc->ds.max_cell;
        out->kernel_launches = launches;
        return PS_OK;
    }
};

// Several ranks of one scene emulated in ONE process on ONE GPU: the same per-rank stages, with device-to-device
// copies where the multi-process path has NCCL messages, driven in rank order (so the hand-over chain of a
// sub-step simply runs left to right).  This is how the slab protocol is tested on a single-GPU box.
template <class Real>
struct SlabGroupLocal {
    std::vector<SlabCollider<Real>*> ranks;
    Error err;

    ~SlabGroupLocal() { for (auto* r : ranks) { r->destroy(); delete r; } }

    int create(const ps_config& cfg, int nranks, uint32_t rows, uint32_t cols) {
        for (int r = 0; r < nranks; ++r) {
            auto* c = new SlabCollider<Real>();
            ranks.push_back(c);
            const int rc = c->create(cfg, r, nranks, nullptr, rows, cols);
            if (rc != PS_OK) { err = c->err; return rc; }
        }
        return PS_OK;
    }

    int copy(void* dst, SlabCollider<Real>* to, const void* src, SlabCollider<Real>* from) {
        // the source was written on `from`'s stream: wait for it, then copy on the receiver's stream
        PS_CUDA_TRY(err, cudaStreamSynchronize(from->stream));
        PS_CUDA_TRY(err, cudaMemcpyAsync(dst, src, from->halo_msg_bytes(), cudaMemcpyDeviceToDevice, to->stream));
        return PS_OK;
    }

    int substep() {
        const int n = (int)ranks.size();
        int rc;
        for (auto* r : ranks) if ((rc = r->stage_bin())) { err = r->err; return rc; }
        for (int k = 0; k < n; ++k) {                                               // M0
            if (k > 0 && (rc = copy(ranks[k - 1]->recv_r, ranks[k - 1], ranks[k]->send_l, ranks[k]))) return rc;
            if (k + 1 < n && (rc = copy(ranks[k + 1]->recv_l, ranks[k + 1], ranks[k]->send_r, ranks[k]))) return rc;
        }
        for (auto* r : ranks) if ((rc = r->stage_build())) { err = r->err; return rc; }
        // M1 chain, left to right, in the order the multi-process path queues it on each rank: a rank whose right boundary was
        // complete after phase A packs its hand-over BEFORE its own phase B (the branch every interior rank of a production
        // run takes), otherwise after it.
        for (int k = 0; k < n; ++k) {
            auto* r = ranks[k];
            const bool early = k + 1 < n && r->send_now_step;
            if (early && (rc = r->pack_m1())) { err = r->err; return rc; }
            if (k > 0 && (rc = r->stage_phase_b())) { err = r->err; return rc; }
            if (k + 1 < n) {
                if (!early && (rc = r->pack_m1())) { err = r->err; return rc; }
                if ((rc = copy(ranks[k + 1]->recv_l, ranks[k + 1], r->send_r, r))) return rc;
            }
        }
        for (auto* r : ranks) if ((rc = r->stage_finish())) { err = r->err; return rc; }
        return PS_OK;
    }
};

}  // namespace ps
