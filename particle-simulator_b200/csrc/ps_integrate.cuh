// ps_integrate.cuh — the integrate loop (src/core.zig:248-260) fused with _applyBoundary
// (src/core.zig:290-318): one streaming pass over the SoA state per sub-step.
//
// HBM-bound: 4 reads + 4 writes of one Real per particle-update (32 B fp32 / 64 B f64) and
// ~12 flops.  Every access is a 128-bit, fully coalesced load or store (float4 / double2):
// each thread owns 4 consecutive particles (fp32) so a warp moves 512 contiguous bytes per
// array per instruction, and all loads of a thread are issued before the first use.
// Compiled with -fmad=false: every multiply and add below is rounded separately, as Zig's
// strict float mode does.
#pragma once
#include "ps_common.cuh"

namespace ps {

// One particle component through one sub-step.  a_dt = acc*dt (already rounded), so
//   vel = vel + acc*dt ; pos = pos + vel*dt          src/core.zig:258-259
//   m = pos<lo | pos>hi ; vel = m ? vel*neg : vel     :306-313
//   pos = min(max(pos, lo), hi)                       :316
template <class Real>
__device__ __forceinline__ void step_component(Real& x, Real& v, Real a_dt, Real dt, Real lo, Real hi, Real neg) {
    v = v + a_dt;
    x = x + v * dt;
    const bool m = (x < lo) | (x > hi);
    v = m ? v * neg : v;
    const Real t = (x > lo) ? x : lo;
    x = (t < hi) ? t : hi;
}

template <class Real> struct Vec;   // 16-byte vector of Real
template <> struct Vec<float>  { using type = float4;  static constexpr int N = 4; };
template <> struct Vec<double> { using type = double2; static constexpr int N = 2; };

__device__ __forceinline__ float&  lane(float4& v, int i)  { return (&v.x)[i]; }
__device__ __forceinline__ double& lane(double2& v, int i) { return (&v.x)[i]; }

// Streaming (evict-first) 128-bit accesses: at 64M particles the state is 1-2 GiB, far
// beyond the 126 MB L2, so nothing read here is worth keeping in cache.
template <class V> __device__ __forceinline__ V ld_stream(const V* p) { return __ldcs(p); }
template <class V> __device__ __forceinline__ void st_stream(V* p, const V& v) { __stcs(p, v); }

// nvec = number of 16-byte vectors per array (arrays are padded to a multiple of 4 elements).
// VPT vectors per thread, block-strided so that each warp instruction stays coalesced.
// SUBSTEPS_FUSED > 1 keeps the particle in registers for several sub-steps (integration-only
// scenes; particles are independent so the result is bit-identical to separate passes).
template <class Real, int VPT, bool STREAM>
__global__ void __launch_bounds__(256)
integrate_bounce_kernel(Real* __restrict__ px, Real* __restrict__ py, Real* __restrict__ vx, Real* __restrict__ vy,
                        uint64_t nvec, StepConst<Real> k, int fused) {
    using V = typename Vec<Real>::type;
    constexpr int N = Vec<Real>::N;
    V* __restrict__ PX = reinterpret_cast<V*>(px);
    V* __restrict__ PY = reinterpret_cast<V*>(py);
    V* __restrict__ VX = reinterpret_cast<V*>(vx);
    V* __restrict__ VY = reinterpret_cast<V*>(vy);
    const uint64_t base = (uint64_t)blockIdx.x * (blockDim.x * VPT) + threadIdx.x;

    V x[VPT], y[VPT], u[VPT], w[VPT];
#pragma unroll
    for (int j = 0; j < VPT; ++j) {
        const uint64_t i = base + (uint64_t)j * blockDim.x;
        if (i < nvec) {
            if (STREAM) { x[j] = ld_stream(PX + i); y[j] = ld_stream(PY + i); u[j] = ld_stream(VX + i); w[j] = ld_stream(VY + i); }
            else        { x[j] = PX[i]; y[j] = PY[i]; u[j] = VX[i]; w[j] = VY[i]; }
        }
    }
    for (int s = 0; s < fused; ++s) {
#pragma unroll
        for (int j = 0; j < VPT; ++j) {
#pragma unroll
            for (int c = 0; c < N; ++c) {
                step_component<Real>(lane(x[j], c), lane(u[j], c), k.ax_dt, k.dt, k.lo, k.hi_x, k.neg_damp);
                step_component<Real>(lane(y[j], c), lane(w[j], c), k.ay_dt, k.dt, k.lo, k.hi_y, k.neg_damp);
            }
        }
    }
#pragma unroll
    for (int j = 0; j < VPT; ++j) {
        const uint64_t i = base + (uint64_t)j * blockDim.x;
        if (i < nvec) {
            if (STREAM) { st_stream(PX + i, x[j]); st_stream(PY + i, y[j]); st_stream(VX + i, u[j]); st_stream(VY + i, w[j]); }
            else        { PX[i] = x[j]; PY[i] = y[j]; VX[i] = u[j]; VY[i] = w[j]; }
        }
    }
}

// Launch `fused` sub-steps in one pass over n_padded elements per array.
// variant: 0 = auto; otherwise VPT*2 + STREAM encoded as {1:VPT1 plain, 2:VPT1 stream, 3:VPT2 plain,
// 4:VPT2 stream, 5:VPT4 plain, 6:VPT4 stream} (tuning hook, see bench.py --sweep).
template <class Real>
inline cudaError_t launch_integrate_bounce(Real* px, Real* py, Real* vx, Real* vy, uint64_t n_padded,
                                           const StepConst<Real>& k, int fused, int variant, cudaStream_t st) {
    constexpr int N = Vec<Real>::N;
    const uint64_t nvec = n_padded / N;
    if (nvec == 0) return cudaSuccess;
    if (variant == 0) {
        // State that fits comfortably in L2 is re-read next sub-step: keep it cached.
        const uint64_t bytes = n_padded * sizeof(Real) * 4;
        const bool big = bytes > (uint64_t)96 << 20;
        variant = sizeof(Real) == 4 ? (big ? 4 : 3) : (big ? 4 : 3);
    }
    const int vpt = variant <= 2 ? 1 : (variant <= 4 ? 2 : 4);
    const bool stream = (variant % 2) == 0;
    const unsigned block = 256;
    const uint64_t per_block = (uint64_t)block * vpt;
    const uint64_t grid = (nvec + per_block - 1) / per_block;
    if (grid > 0x7fffffffull) return cudaErrorInvalidConfiguration;
#define PS_LAUNCH(VPT_, STREAM_) \
    integrate_bounce_kernel<Real, VPT_, STREAM_><<<(unsigned)grid, block, 0, st>>>(px, py, vx, vy, nvec, k, fused)
    if (vpt == 1) { if (stream) PS_LAUNCH(1, true); else PS_LAUNCH(1, false); }
    else if (vpt == 2) { if (stream) PS_LAUNCH(2, true); else PS_LAUNCH(2, false); }
    else { if (stream) PS_LAUNCH(4, true); else PS_LAUNCH(4, false); }
#undef PS_LAUNCH
    return cudaGetLastError();
}

}  // namespace ps
