// oracle/oracle_capi.cpp — plain-C entry points over core_ref.hpp so that tests/ and
// bench.py's CPU-baseline legs can drive the oracle through ctypes.
// TEST INFRASTRUCTURE ONLY (see the header of core_ref.hpp).  PARITY UNPINNED.
//
// Build: g++ -O2 -ffp-contract=off -fPIC -shared (oracle/Makefile).
#include "core_ref.hpp"

#include <chrono>
#include <cstring>

namespace {

template <class Real>
int run(const oracle::Params* p, uint64_t n, Real* pos, Real* vel, Real* acc,
        uint64_t substeps, int collisions, int t_acc,
        int64_t capture_substep, uint64_t* pairs_out, uint64_t pairs_cap, uint64_t* pairs_count,
        oracle::Stats* stats, double* seconds) {
    if (!p || (n && (!pos || !vel || !acc))) return -1;
    if ((int32_t)p->grid_size <= 0) return -2;
    oracle::Sim<Real> sim;
    sim.init(*p, (size_t)n, pos, vel, acc);
    sim.t_acc = t_acc != 0;
    if (collisions && (sim.rows == 0 || sim.cols == 0)) return -2;
    if (pairs_count) *pairs_count = 0;
    const auto t0 = std::chrono::steady_clock::now();
    for (uint64_t s = 0; s < substeps; ++s) {
        sim.substep(collisions != 0);
        if (collisions && (int64_t)s == capture_substep && pairs_count) {
            const uint64_t np = sim.collision_pairs.size() / 2;
            *pairs_count = np;
            if (pairs_out) {
                const uint64_t m = np < pairs_cap ? np : pairs_cap;
                std::memcpy(pairs_out, sim.collision_pairs.data(), (size_t)m * 2 * sizeof(uint64_t));
            }
        }
    }
    const auto t1 = std::chrono::steady_clock::now();
    if (seconds) *seconds = std::chrono::duration<double>(t1 - t0).count();
    if (stats) *stats = sim.stats;
    return 0;
}

}  // namespace

extern "C" {

// Advance `substeps` iterations of the sub-step loop (src/core.zig:247-267) in place.
// pos/vel/acc are interleaved (x,y) arrays of length 2n, exactly []@Vector(2,f64).
// capture_substep: index of the sub-step whose ordered pair list is copied to
// pairs_out ({i, j} per entry, reference order); -1 for none.
int oracle_run_f64(const oracle::Params* p, uint64_t n, double* pos, double* vel, double* acc,
                   uint64_t substeps, int collisions, int t_acc,
                   int64_t capture_substep, uint64_t* pairs_out, uint64_t pairs_cap, uint64_t* pairs_count,
                   oracle::Stats* stats, double* seconds) {
    return run<double>(p, n, pos, vel, acc, substeps, collisions, t_acc, capture_substep,
                       pairs_out, pairs_cap, pairs_count, stats, seconds);
}

int oracle_run_f32(const oracle::Params* p, uint64_t n, float* pos, float* vel, float* acc,
                   uint64_t substeps, int collisions, int t_acc,
                   int64_t capture_substep, uint64_t* pairs_out, uint64_t pairs_cap, uint64_t* pairs_count,
                   oracle::Stats* stats, double* seconds) {
    return run<float>(p, n, pos, vel, acc, substeps, collisions, t_acc, capture_substep,
                      pairs_out, pairs_cap, pairs_count, stats, seconds);
}

// Grid shape as SpatialHashGrid.init derives it (src/core.zig:61-67).
int oracle_grid_shape(const oracle::Params* p, uint64_t* rows, uint64_t* cols) {
    if (!p || (int32_t)p->grid_size <= 0) return -2;
    oracle::Sim<double> sim;
    double dummy = 0;
    sim.init(*p, 0, &dummy, &dummy, &dummy);
    *rows = sim.rows; *cols = sim.cols;
    return 0;
}

// Derived dt as Solver.init computes it in each scalar type (src/core.zig:227-228).
double oracle_dt_f64(const oracle::Params* p) {
    return (1.0 / p->target_fps) / (double)p->sub_steps;
}
float oracle_dt_f32(const oracle::Params* p) {
    return ((float)1 / (float)p->target_fps) / (float)p->sub_steps;
}

}  // extern "C"
