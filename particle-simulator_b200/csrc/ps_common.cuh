// ps_common.cuh — shared definitions for libparticles_cuda (sm_100a only).
#pragma once
#include <cuda_runtime.h>

#include <cmath>
#include <cstdint>
#include <cstdio>
#include <string>

#include "particles_cuda.h"

namespace ps {

// Scalar constants of one sub-step, derived on the host in the working type exactly as
// Solver.init / _applyBoundary derive them (src/core.zig:227-228, :301-303) and passed to
// kernels by value.  Everything is a run-time value so that nvcc cannot fold `v + 0*dt`
// (which turns -0.0 into +0.0 and must be kept).
template <class Real>
struct StepConst {
    Real dt;          // Solver.dt                                   src/core.zig:228
    Real ax_dt;       // zeroNearZero(acc.x) * dt  (acc.x is always 0)   :254-258
    Real ay_dt;       // zeroNearZero(t_acc ? g : 0) * dt
    Real lo;          // min_bound = radius                          :301
    Real hi_x, hi_y;  // max_bound = (width - radius, height - radius)   :302
    Real neg_damp;    // -boundary_damping                           :303
    // collision constants
    Real radius;
    Real r2sq;        // (2r)*(2r)                                   :341-342
    Real min_dis;     // 2.0 * radius                                :372
    Real k_coef;      // -(1.0 + particle_damping)                   :368
    Real grid_size;   // SpatialHashGrid.grid_size                   :75
    Real inv_gs;      // 1 / grid_size, used only when grid_size is a power of two (then x * inv_gs == x / grid_size exactly)
    int gs_pow2;
    uint32_t rows, cols;
    int t_acc;
    uint32_t keep_from;   // resolve: slots from here on get their head word written even when their sub-list has ended (slab handles: the
                          // two boundary columns that slab_right_check_kernel inspects); 0xffffffff on a single GPU.  Set inside the kernel.
};

template <class Real>
inline StepConst<Real> make_step_const(const ps_config& c, bool t_acc, uint32_t rows, uint32_t cols) {
    StepConst<Real> k{};
    const Real fps = (Real)c.target_fps;
    const Real delta_time = (Real)1 / fps;                      // :227
    k.dt = delta_time / (Real)c.sub_steps;                      // :228
    const Real ax = (Real)0;
    Real ay = t_acc ? (Real)c.g : (Real)0;                      // :249-253
    if ((ay < 0 ? -ay : ay) < (Real)1e-6) ay = (Real)0;         // :242-244
    // volatile keeps the host compiler from contracting or widening these products
    volatile Real axdt = ax * k.dt, aydt = ay * k.dt;
    k.ax_dt = axdt; k.ay_dt = aydt;
    const Real r = (Real)c.radius;
    k.lo = r;
    volatile Real hx = (Real)c.width - r, hy = (Real)c.height - r;
    k.hi_x = hx; k.hi_y = hy;
    k.neg_damp = -(Real)c.boundary_damping;
    k.radius = r;
    volatile Real r2 = (Real)2.0 * r;
    volatile Real r2sq = r2 * r2;
    k.r2sq = r2sq;
    k.min_dis = r2;
    k.k_coef = -((Real)1.0 + (Real)c.particle_damping);
    k.grid_size = (Real)c.grid_size;
    {
        int e = 0;
        const double m = std::frexp((double)k.grid_size, &e);       // grid_size = m * 2^e, m in [0.5, 1)
        k.gs_pow2 = (k.grid_size > (Real)0 && m == 0.5 && e > -100 && e < 100) ? 1 : 0;
        k.inv_gs = k.gs_pow2 ? (Real)std::ldexp(1.0, 1 - e) : (Real)0;
    }
    k.rows = rows; k.cols = cols;
    k.t_acc = t_acc ? 1 : 0;
    return k;
}

struct Error {
    int status = PS_OK;
    std::string text;
};

#define PS_CUDA_TRY(sim_err, expr)                                                          \
    do {                                                                                    \
        cudaError_t e__ = (expr);                                                           \
        if (e__ != cudaSuccess) {                                                           \
            (sim_err).status = (e__ == cudaErrorMemoryAllocation) ? PS_ERR_OOM : PS_ERR_CUDA; \
            (sim_err).text = std::string(#expr) + ": " + cudaGetErrorString(e__);           \
            return (sim_err).status;                                                        \
        }                                                                                   \
    } while (0)

inline uint64_t round_up(uint64_t x, uint64_t m) { return (x + m - 1) / m * m; }

}  // namespace ps
