// ps_scan.cuh — exclusive prefix sum over u8 or u32 items (computeStartArray, src/core.zig:137-142, over the u8 cell
// counters; the pair-capture offsets over u32).  Three launches: per-tile reduce, a one-block scan of the tile sums,
// per-tile scan + offset.  HBM traffic per element: the input read twice, a u32 written = 6 B for the u8 counters,
// 12 B for u32 items.
#pragma once
#include "ps_common.cuh"

namespace ps {

constexpr int SCAN_THREADS = 256;
constexpr int SCAN_ITEMS = 16;                          // per thread, as 4 x uint4
constexpr int SCAN_TILE = SCAN_THREADS * SCAN_ITEMS;    // 4096 elements per CTA

__device__ __forceinline__ uint32_t warp_inclusive_scan(uint32_t v) {
    const unsigned lane = threadIdx.x & 31;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const uint32_t t = __shfl_up_sync(0xffffffffu, v, d);
        if (lane >= (unsigned)d) v += t;
    }
    return v;
}

// Block-wide exclusive scan of one value per thread; returns the exclusive prefix, *total = block sum.
template <int THREADS>
__device__ __forceinline__ uint32_t block_exclusive_scan(uint32_t v, uint32_t* total) {
    __shared__ uint32_t warp_sums[THREADS / 32];
    const unsigned lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const uint32_t inc = warp_inclusive_scan(v);
    if (lane == 31) warp_sums[wid] = inc;
    __syncthreads();
    if (wid == 0) {
        uint32_t s = lane < THREADS / 32 ? warp_sums[lane] : 0;
        s = warp_inclusive_scan(s);
        if (lane < THREADS / 32) warp_sums[lane] = s;
    }
    __syncthreads();
    const uint32_t base = wid ? warp_sums[wid - 1] : 0;
    *total = warp_sums[THREADS / 32 - 1];
    __syncthreads();
    return base + inc - v;
}

__device__ __forceinline__ void load_items(const uint32_t* __restrict__ in, uint64_t len, uint64_t start, uint32_t (&x)[SCAN_ITEMS]) {
    if (start + SCAN_ITEMS <= len && ((reinterpret_cast<uintptr_t>(in + start) & 15) == 0)) {
        const uint4* p = reinterpret_cast<const uint4*>(in + start);
#pragma unroll
        for (int j = 0; j < SCAN_ITEMS / 4; ++j) {
            const uint4 q = p[j];
            x[4 * j] = q.x; x[4 * j + 1] = q.y; x[4 * j + 2] = q.z; x[4 * j + 3] = q.w;
        }
    } else {
#pragma unroll
        for (int j = 0; j < SCAN_ITEMS; ++j) x[j] = (start + j < len) ? in[start + j] : 0u;
    }
}

// Same for one-byte items (the per-cell counters are u8 like the reference's grid_array, src/core.zig:51, four to a
// 32-bit word): 16 items = one 16-byte load.  The buffer is padded to a whole number of 16-byte groups with zeros.
__device__ __forceinline__ void load_items(const uint8_t* __restrict__ in, uint64_t len, uint64_t start, uint32_t (&x)[SCAN_ITEMS]) {
    if (start < len) {
        const uint4 q = *reinterpret_cast<const uint4*>(in + start);
        const uint32_t w[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
        for (int j = 0; j < SCAN_ITEMS; ++j) x[j] = (w[j >> 2] >> ((j & 3) * 8)) & 0xffu;
    } else {
#pragma unroll
        for (int j = 0; j < SCAN_ITEMS; ++j) x[j] = 0u;
    }
}

// sums[b] = sum of tile b; optionally track the largest element (max cell population).
template <class In>
__global__ void __launch_bounds__(SCAN_THREADS)
scan_reduce_kernel(const In* __restrict__ in, uint64_t len, uint32_t* __restrict__ sums, uint32_t* __restrict__ max_out) {
    const uint64_t start = (uint64_t)blockIdx.x * SCAN_TILE + (uint64_t)threadIdx.x * SCAN_ITEMS;
    uint32_t x[SCAN_ITEMS];
    load_items(in, len, start, x);
    uint32_t s = 0, m = 0;
#pragma unroll
    for (int j = 0; j < SCAN_ITEMS; ++j) { s += x[j]; m = x[j] > m ? x[j] : m; }
    __shared__ uint32_t sh_s[SCAN_THREADS / 32], sh_m[SCAN_THREADS / 32];
#pragma unroll
    for (int d = 16; d; d >>= 1) {
        s += __shfl_xor_sync(0xffffffffu, s, d);
        const uint32_t o = __shfl_xor_sync(0xffffffffu, m, d);
        m = o > m ? o : m;
    }
    if ((threadIdx.x & 31) == 0) { sh_s[threadIdx.x >> 5] = s; sh_m[threadIdx.x >> 5] = m; }
    __syncthreads();
    if (threadIdx.x == 0) {
        uint32_t ts = 0, tm = 0;
#pragma unroll
        for (int w = 0; w < SCAN_THREADS / 32; ++w) { ts += sh_s[w]; tm = sh_m[w] > tm ? sh_m[w] : tm; }
        sums[blockIdx.x] = ts;
        if (max_out && tm > 0u) atomicMax(max_out, tm);     // one reduction atomic per CTA at most
    }
}

// In-place exclusive scan of the tile sums by one CTA; writes the grand total to out_total.
__global__ void __launch_bounds__(1024)
scan_spine_kernel(uint32_t* __restrict__ sums, uint32_t nb, uint32_t* __restrict__ out_total) {
    __shared__ uint32_t carry_sh;
    if (threadIdx.x == 0) carry_sh = 0;
    __syncthreads();
    for (uint32_t base = 0; base < nb; base += 1024) {
        const uint32_t i = base + threadIdx.x;
        const uint32_t v = i < nb ? sums[i] : 0u;
        uint32_t total;
        const uint32_t ex = block_exclusive_scan<1024>(v, &total);
        const uint32_t carry = carry_sh;
        if (i < nb) sums[i] = carry + ex;
        __syncthreads();
        if (threadIdx.x == 0) carry_sh = carry + total;
        __syncthreads();
    }
    if (threadIdx.x == 0) *out_total = carry_sh;
}

// out[i] = sums[tile] + exclusive prefix within the tile; out may alias in (32-bit items).
template <class In>
__global__ void __launch_bounds__(SCAN_THREADS)
scan_down_kernel(const In* in, uint64_t len, const uint32_t* __restrict__ sums, uint32_t* out) {
    const uint64_t start = (uint64_t)blockIdx.x * SCAN_TILE + (uint64_t)threadIdx.x * SCAN_ITEMS;
    uint32_t x[SCAN_ITEMS];
    load_items(in, len, start, x);
    uint32_t s = 0;
#pragma unroll
    for (int j = 0; j < SCAN_ITEMS; ++j) s += x[j];
    uint32_t total;
    uint32_t run = block_exclusive_scan<SCAN_THREADS>(s, &total) + sums[blockIdx.x];
    uint32_t y[SCAN_ITEMS];
#pragma unroll
    for (int j = 0; j < SCAN_ITEMS; ++j) { y[j] = run; run += x[j]; }
    if (start + SCAN_ITEMS <= len && ((reinterpret_cast<uintptr_t>(out + start) & 15) == 0)) {
        uint4* p = reinterpret_cast<uint4*>(out + start);
#pragma unroll
        for (int j = 0; j < SCAN_ITEMS / 4; ++j) p[j] = make_uint4(y[4 * j], y[4 * j + 1], y[4 * j + 2], y[4 * j + 3]);
    } else {
#pragma unroll
        for (int j = 0; j < SCAN_ITEMS; ++j) if (start + j < len) out[start + j] = y[j];
    }
}

inline uint64_t scan_tiles(uint64_t len) { return (len + SCAN_TILE - 1) / SCAN_TILE; }

// out[0..len) = exclusive scan of in[0..len); *total_out (device) = sum.  `sums` holds scan_tiles(len) words.
template <class In>
inline cudaError_t exclusive_scan(const In* in, uint32_t* out, uint64_t len, uint32_t* sums,
                                  uint32_t* total_out, uint32_t* max_out, cudaStream_t st) {
    if (len == 0) return cudaMemsetAsync(total_out, 0, sizeof(uint32_t), st);
    const uint64_t nb = scan_tiles(len);
    scan_reduce_kernel<In><<<(unsigned)nb, SCAN_THREADS, 0, st>>>(in, len, sums, max_out);
    scan_spine_kernel<<<1, 1024, 0, st>>>(sums, (uint32_t)nb, total_out);
    scan_down_kernel<In><<<(unsigned)nb, SCAN_THREADS, 0, st>>>(in, len, sums, out);
    return cudaGetLastError();
}
inline cudaError_t exclusive_scan_u32(const uint32_t* in, uint32_t* out, uint64_t len, uint32_t* sums,
                                      uint32_t* total_out, uint32_t* max_out, cudaStream_t st) {
    return exclusive_scan<uint32_t>(in, out, len, sums, total_out, max_out, st);
}

// Per-cell counters, one byte each, incremented through the enclosing 32-bit word.  Returns the counter's previous
// value (the slot the caller gets inside its cell).  A counter that would pass 255 carries into its neighbour: the
// caller flags that as PS_ERR_CELL_OVERFLOW (the reference's u8 wraps or panics there, src/core.zig:132).
__device__ __forceinline__ uint32_t cell_count_fetch_inc(uint32_t* __restrict__ count_words, uint32_t c) {
    const uint32_t sh = (c & 3u) * 8u;
    return (atomicAdd(count_words + (c >> 2), 1u << sh) >> sh) & 0xffu;
}
inline uint64_t cell_count_bytes(uint64_t ncell) { return (ncell + 15) / 16 * 16; }   // whole 16-byte groups

}  // namespace ps
