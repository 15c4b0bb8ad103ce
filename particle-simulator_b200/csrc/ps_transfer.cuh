// ps_transfer.cuh — host <-> device state movement.
// The host side keeps core.zig's layout ([]@Vector(2,f64): interleaved x,y; src/core.zig:5,14-16);
// the device keeps a structure of arrays in the working type.  These kernels do the
// (de)interleave and the f64<->f32 conversion in one pass.  When the host buffer is pinned
// (ps_host_alloc / cudaHostRegister) they read or write it directly over PCIe (no staging
// copy); otherwise the data goes through a device staging buffer.
#pragma once
#include "ps_common.cuh"

namespace ps {

template <class T> struct Pair;
template <> struct Pair<double> { using type = double2; };
template <> struct Pair<float>  { using type = float2; };

// src: n interleaved (x,y) of type H  ->  ax[i], ay[i] of type D
template <class H, class D>
__global__ void __launch_bounds__(256)
deinterleave_kernel(const typename Pair<H>::type* __restrict__ src, D* __restrict__ ax, D* __restrict__ ay, uint64_t n) {
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        const typename Pair<H>::type p = src[i];
        ax[i] = (D)p.x;
        ay[i] = (D)p.y;
    }
}

// ax[i], ay[i] of type D  ->  dst: n interleaved (x,y) of type H
template <class D, class H>
__global__ void __launch_bounds__(256)
interleave_kernel(const D* __restrict__ ax, const D* __restrict__ ay, typename Pair<H>::type* __restrict__ dst, uint64_t n) {
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        typename Pair<H>::type p;
        p.x = (H)ax[i];
        p.y = (H)ay[i];
        dst[i] = p;
    }
}

inline unsigned transfer_grid(uint64_t n) {
    const uint64_t want = (n + 255) / 256;
    const uint64_t cap = 148ull * 16;   // 16 resident CTAs of 256 threads per SM, grid-stride beyond
    return (unsigned)(want < cap ? (want ? want : 1) : cap);
}

// Is `p` page-locked host memory that kernels on the current device can address directly?
inline bool host_ptr_is_pinned(const void* p, void** dev_alias) {
    cudaPointerAttributes a{};
    if (cudaPointerGetAttributes(&a, p) != cudaSuccess) { cudaGetLastError(); return false; }
    if (a.type == cudaMemoryTypeHost && a.devicePointer) { *dev_alias = a.devicePointer; return true; }
    return false;
}

}  // namespace ps
