// oracle/core_ref.hpp — CPU restatement of the reference particle step.
//
// TEST INFRASTRUCTURE ONLY.  Nothing shipped in libparticles_cuda may include,
// link or call this file; only tests/, __graft_entry__.smoke() and bench.py's
// cpu_baseline / --impl reference legs use it, as the checker or the timed
// CPU baseline.
//
// PARITY UNPINNED: the reference (Zig 0.14, /root/reference/src/core.zig) has no
// tests, golden vectors or fixtures, and no Zig toolchain exists in this image,
// so this restatement cannot be checked against reference output.  It is pinned
// only by (a) hand-derived known-answer tests written from core.zig's formulas
// (tests/test_oracle_kat.py) and (b) an independent second restatement in pure
// Python (oracle/core_ref.py) that must agree bit-for-bit at f64.
//
// Everything follows src/core.zig operation for operation; each function cites
// the lines it restates.  Build with -O2 -ffp-contract=off (Zig's default float
// mode is strict: no reassociation, no implicit fused multiply-add).
//
// Real = double  : "core.zig as written" (Vec2f64, src/core.zig:5).
// Real = float   : the same operation order with every f64 replaced by f32
//                  (the north_star's fp32 device layout); parameters are cast to
//                  float first and derived quantities (dt, bounds) are computed
//                  in float.
#pragma once
#include <cmath>
#include <cstddef>
#include <cstdint>
#include <vector>

namespace oracle {

// Scene parameters as main.zig passes them (src/main.zig:25-29,63-75).
struct Params {
    double radius;             // Particles.init radius              (src/main.zig:64)
    double width, height;      // updatePhysics(width, height)       (src/main.zig:109)
    double grid_size;          // SpatialHashGrid.init grid_size     (src/main.zig:74)
    double target_fps;         // Solver.init target_fps             (src/main.zig:72)
    uint64_t sub_steps;        // Solver.init sub_steps
    double g;                  // Solver.init g
    double boundary_damping;   // Solver.init boundary_damping
    double particle_damping;   // Solver.init particle_damping
};

struct Stats {
    uint64_t pairs_total;      // sum over substeps of collision_pairs.items.len
    uint64_t pairs_last;       // list length of the last substep run
    uint32_t max_cell_count;   // largest per-cell population seen (reference counter is u8, src/core.zig:51)
    uint32_t cell_overflow;    // 1 if any cell held > 255 particles (reference would wrap/panic, src/core.zig:132)
};

template <class Real>
struct Sim {
    // ---- Particles (src/core.zig:13-39): interleaved (x,y) like []@Vector(2,f64)
    size_t n = 0;
    Real radius = 0;
    Real* pos = nullptr;   // 2n
    Real* vel = nullptr;   // 2n
    Real* acc = nullptr;   // 2n

    // ---- Solver (src/core.zig:217-238)
    Real delta_time = 0, dt = 0, g = 0, boundary_damping = 0, particle_damping = 0;
    size_t sub_steps = 0;
    bool t_acc = false;            // src/core.zig:220

    // ---- SpatialHashGrid (src/core.zig:47-85)
    size_t rows = 0, cols = 0;
    Real grid_size = 0;
    std::vector<uint32_t> grid_array;    // u8 in the reference; 32-bit here + overflow flag
    std::vector<size_t> start_array, cell_counters, sorted_array;
    std::vector<uint64_t> collision_pairs;   // flattened [2]usize entries: {particle_i, particle_j}

    Real width = 0, height = 0;
    Stats stats{};

    // Solver.init (src/core.zig:226-238) + SpatialHashGrid.init (src/core.zig:60-85)
    void init(const Params& p, size_t count, Real* pos_, Real* vel_, Real* acc_) {
        n = count; pos = pos_; vel = vel_; acc = acc_;
        radius = (Real)p.radius;
        width = (Real)p.width; height = (Real)p.height;
        const Real fps = (Real)p.target_fps;
        delta_time = (Real)1 / fps;                                   // :227
        dt = delta_time / (Real)p.sub_steps;                          // :228
        sub_steps = (size_t)p.sub_steps;
        g = (Real)p.g;
        boundary_damping = (Real)p.boundary_damping;
        particle_damping = (Real)p.particle_damping;
        grid_size = (Real)p.grid_size;
        const int32_t wi = (int32_t)p.width, hi = (int32_t)p.height, gi = (int32_t)p.grid_size;  // :61-63 (@intFromFloat truncates)
        cols = (size_t)floordiv(wi, gi);                              // :66
        rows = (size_t)floordiv(hi, gi);                              // :67
        grid_array.assign(rows * cols, 0);
        start_array.assign(rows * cols, 0);
        cell_counters.assign(rows * cols, 0);
        sorted_array.assign(n, 0);
        collision_pairs.clear();
        stats = Stats{};
    }

    static int32_t floordiv(int32_t a, int32_t b) {                   // @divFloor on i32
        int32_t q = a / b, r = a % b;
        return (r != 0 && ((r < 0) != (b < 0))) ? q - 1 : q;
    }

    static Real zeroNearZero(Real x) {                                // :242-244
        return (std::fabs(x) < (Real)1e-6) ? (Real)0 : x;
    }

    // integrate loop, src/core.zig:248-260
    void integrate() {
        for (size_t i = 0; i < n; ++i) {
            Real* a = acc + 2 * i; Real* v = vel + 2 * i; Real* x = pos + 2 * i;
            if (t_acc) { a[1] = g; } else { a[0] = (Real)0; a[1] = (Real)0; }   // :249-253
            a[0] = zeroNearZero(a[0]); a[1] = zeroNearZero(a[1]);               // :254-257
            v[0] = v[0] + a[0] * dt; v[1] = v[1] + a[1] * dt;                   // :258
            x[0] = x[0] + v[0] * dt; x[1] = x[1] + v[1] * dt;                   // :259
        }
    }

    // _applyBoundary, src/core.zig:290-318 (the variant updatePhysics calls, :261)
    void applyBoundary() {
        const Real min_b[2] = {radius, radius};                                 // :301
        const Real max_b[2] = {width - radius, height - radius};                // :302
        const Real neg = -boundary_damping;                                     // :303
        for (size_t i = 0; i < n; ++i) {
            for (int c = 0; c < 2; ++c) {
                Real& x = pos[2 * i + c]; Real& v = vel[2 * i + c];
                const bool m = (x < min_b[c]) | (x > max_b[c]);                 // :306-310
                v = m ? v * neg : v;                                            // :313
                Real t = (x > min_b[c]) ? x : min_b[c];                         // @max(pos, min)  :316
                x = (t < max_b[c]) ? t : max_b[c];                              // @min(.., max)
            }
        }
    }

    // clear, src/core.zig:95-100
    void clear() {
        std::fill(grid_array.begin(), grid_array.end(), 0u);
        std::fill(cell_counters.begin(), cell_counters.end(), (size_t)0);
        std::fill(start_array.begin(), start_array.end(), (size_t)0);
        std::fill(sorted_array.begin(), sorted_array.end(), (size_t)0);
    }

    // computeCoordinate + computeGridIndex, src/core.zig:103-121
    size_t cellOf(Real x, Real y) const {
        size_t i = (size_t)std::floor(x / grid_size);                           // :104
        size_t j = (size_t)std::floor(y / grid_size);                           // :105
        if (i > cols - 1) i = cols - 1;                                         // :108
        if (j > rows - 1) j = rows - 1;                                         // :109
        return i * rows + j;                                                    // :120 (column-major)
    }

    // computeGridArray, src/core.zig:124-134
    void computeGridArray() {
        for (size_t i = 0; i < n; ++i) {
            const size_t c = cellOf(pos[2 * i], pos[2 * i + 1]);
            const uint32_t k = ++grid_array[c];                                 // :132
            if (k > stats.max_cell_count) stats.max_cell_count = k;
            if (k > 255u) stats.cell_overflow = 1;
        }
    }

    // computeStartArray, src/core.zig:137-142
    void computeStartArray() {
        if (grid_array.empty()) return;
        start_array[0] = 0;
        for (size_t i = 1; i < grid_array.size(); ++i)
            start_array[i] = start_array[i - 1] + grid_array[i - 1];
    }

    // computeSortedArray, src/core.zig:145-159
    void computeSortedArray() {
        for (size_t i = 0; i < n; ++i) {
            const size_t c = cellOf(pos[2 * i], pos[2 * i + 1]);
            sorted_array[start_array[c] + cell_counters[c]] = i;                // :156
            cell_counters[c] += 1;                                              // :157
        }
    }

    // distanceSquared, src/core.zig:333-338
    static Real distanceSquared(Real xi, Real yi, Real xj, Real yj) {
        const Real dx_sq = (xj - xi) * (xj - xi);
        const Real dy_sq = (yj - yi) * (yj - yi);
        return dx_sq + dy_sq;
    }
    // checkCollision, src/core.zig:339-346
    bool checkCollision(size_t i, size_t j) const {
        const Real d2 = distanceSquared(pos[2 * i], pos[2 * i + 1], pos[2 * j], pos[2 * j + 1]);
        const Real r_2 = (Real)2.0 * radius;
        return d2 < r_2 * r_2;
    }

    // detect half of detect_collisions, src/core.zig:161-209
    void detect() {
        static const int32_t off[5][2] = {{0, 0}, {0, 1}, {1, 0}, {1, 1}, {-1, 1}};   // :58
        collision_pairs.clear();                                                // :163
        const size_t ncell = cols * rows;
        for (size_t x = 0; x < cols; ++x) {
            for (size_t y = 0; y < rows; ++y) {
                const size_t c = x * rows + y;
                const size_t s = start_array[c];
                const size_t e = (c + 1 < ncell) ? start_array[c + 1] : n;      // :168
                if (s >= e) continue;
                const int32_t gx = (int32_t)x, gy = (int32_t)y;
                for (size_t j = s; j < e; ++j) {
                    const size_t pj = sorted_array[j];
                    for (int o = 0; o < 5; ++o) {
                        const int32_t nx = gx + off[o][0], ny = gy + off[o][1];
                        if (nx < 0 || nx >= (int32_t)cols || ny < 0 || ny >= (int32_t)rows) continue;   // :188-189
                        const size_t nc = (size_t)nx * rows + (size_t)ny;
                        const size_t s2 = start_array[nc];
                        const size_t e2 = (nc + 1 < ncell) ? start_array[nc + 1] : n;
                        for (size_t i = s2; i < e2; ++i) {
                            const size_t pi = sorted_array[i];
                            if (pi == pj) continue;                             // :201
                            if (checkCollision(pi, pj)) {                       // :202
                                collision_pairs.push_back(pi);                  // entry = [particle_i, particle_j]  :203
                                collision_pairs.push_back(pj);
                            }
                        }
                    }
                }
            }
        }
    }

    // resolve_collision, src/core.zig:348-392
    void resolve_collision(size_t i, size_t j) {
        const Real pix = pos[2 * i], piy = pos[2 * i + 1];
        const Real pjx = pos[2 * j], pjy = pos[2 * j + 1];
        const Real vix = vel[2 * i], viy = vel[2 * i + 1];
        const Real vjx = vel[2 * j], vjy = vel[2 * j + 1];
        const Real dist = std::sqrt(distanceSquared(pix, piy, pjx, pjy));       // :354
        if (dist < (Real)0.0000000001) return;                                  // :355
        const Real nx = (pjx - pix) / dist, ny = (pjy - piy) / dist;            // :361
        const Real rvx = vjx - vix, rvy = vjy - viy;                            // :363
        const Real van = rvx * nx + rvy * ny;                                   // :365
        const Real k = -((Real)1.0 + particle_damping) * van / (Real)2;         // :368
        const Real min_dis = (Real)2.0 * radius;                                // :372
        const Real overlap = min_dis - dist;                                    // :374
        const Real corr = overlap * (Real)0.5;                                  // :375
        if (overlap > (Real)0) {                                                // :377
            if (!t_acc) {
                pos[2 * i] -= corr * nx; pos[2 * i + 1] -= corr * ny;           // :379
                pos[2 * j] += corr * nx; pos[2 * j + 1] += corr * ny;           // :380
            } else {
                const Real pax = ((Real)10000.0 * nx) * overlap;                // :385
                const Real pay = ((Real)10000.0 * ny) * overlap;
                vel[2 * i] -= pax * dt; vel[2 * i + 1] -= pay * dt;             // :386
                vel[2 * j] += pax * dt; vel[2 * j + 1] += pay * dt;             // :387
            }
        }
        vel[2 * i] -= nx * k; vel[2 * i + 1] -= ny * k;                         // :390
        vel[2 * j] += nx * k; vel[2 * j + 1] += ny * k;                         // :391
    }

    // resolve half of detect_collisions, src/core.zig:211-213
    void resolveAll() {
        const size_t np = collision_pairs.size() / 2;
        for (size_t k = 0; k < np; ++k)
            resolve_collision((size_t)collision_pairs[2 * k], (size_t)collision_pairs[2 * k + 1]);
        stats.pairs_total += np;
        stats.pairs_last = np;
    }

    // One iteration of the sub-step loop of updatePhysics, src/core.zig:247-267.
    // collisions=false is the "integrate + wall bounce only" mode BASELINE.json's
    // configs name; the reference has no such switch (updatePhysics always runs
    // lines :262-266), so that mode stops after :261.
    void substep(bool collisions) {
        integrate();
        applyBoundary();                                                        // :261
        if (!collisions) return;
        clear();                                                                // :262
        computeGridArray();                                                     // :263
        computeStartArray();                                                    // :264
        computeSortedArray();                                                   // :265
        detect();                                                               // :266
        resolveAll();
    }

    // updatePhysics, src/core.zig:245-268
    void updatePhysics(bool collisions) {
        for (size_t s = 0; s < sub_steps; ++s) substep(collisions);
    }
};

}  // namespace oracle
