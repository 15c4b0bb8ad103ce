// core_cuda.hpp — C++ mirror of src/core.zig's public API over the C ABI (include/particles_cuda.h).
//
// The reference is compiled Zig; no Zig toolchain exists in this image, so the host side above
// the C ABI is written in C++ with the reference's own type and method names (Particles, Solver,
// SpatialHashGrid, Spawner; updatePhysics, toggle_acc, spawnRandom) and the same argument order,
// so that main.zig's init / updateAndRender (src/main.zig:35-76,95-114) can be replayed line for
// line (headless_main.cpp does).  zig/core_cuda.zig is the same shim in Zig, shipped as source.
//
// All arithmetic of the step happens in libparticles_cuda on the GPU; nothing here computes.
#pragma once
#include <cmath>
#include <cstdint>
#include <cstring>
#include <stdexcept>
#include <string>
#include <vector>

#include "particles_cuda.h"

namespace core_cuda {

struct Vec2f64 { double x, y; };               // @Vector(2, f64), src/core.zig:5
struct Color { uint8_t r, g, b, a; };          // sdl.Color, src/core.zig:17

struct Error : std::runtime_error {
    int status;
    Error(int s, const std::string& what) : std::runtime_error(what), status(s) {}
};

inline void check(int rc, const ps_sim* sim, const char* where) {
    if (rc == PS_OK) return;
    const char* text = ps_last_error(sim);
    throw Error(rc, std::string(where) + ": " + ((text && *text) ? text : ps_status_string(rc)));
}

// splitmix64-based counter generator; identical to particle-simulator_b200/scenes.py (uniform01).
inline uint64_t splitmix64(uint64_t z) {
    z += 0x9E3779B97F4A7C15ull;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}
inline double uniform01(uint64_t seed_base, uint64_t index, uint64_t stream) {
    return (double)(splitmix64((index * 4 + stream) ^ seed_base) >> 11) * (1.0 / 9007199254740992.0);
}

// src/colors.zig:6-47
inline Color hsvToSdlColor(float h, float s, float v, uint8_t a) {
    const float c = v * s;
    const float x = c * (1.0f - std::fabs(std::fmod(h / 60.0f, 2.0f) - 1.0f));
    const float m = v - c;
    float r = 0, g = 0, b = 0;
    if (h < 60.0f) { r = c; g = x; }
    else if (h < 120.0f) { r = x; g = c; }
    else if (h < 180.0f) { g = c; b = x; }
    else if (h < 240.0f) { g = x; b = c; }
    else if (h < 300.0f) { r = x; b = c; }
    else { r = c; b = x; }
    return Color{(uint8_t)((r + m) * 255.0f), (uint8_t)((g + m) * 255.0f), (uint8_t)((b + m) * 255.0f), a};
}

// src/core.zig:13-45.  Host mirrors keep the reference's slices; the device handle is bound by Solver.
struct Particles {
    std::vector<Vec2f64> positions, velocities, accelerations;
    std::vector<Color> colors;
    double radius = 0;
    ps_sim* sim = nullptr;      // backend handle (the one thing main.zig gains)
    ps_config bound{};          // configuration `sim` was created with
    bool dirty = true;          // host mirrors newer than the device state

    static Particles init(size_t particles_count, double radius) {
        Particles p;
        p.positions.assign(particles_count, Vec2f64{0, 0});       // :27
        p.velocities.assign(particles_count, Vec2f64{0, 0});      // :28
        p.accelerations.assign(particles_count, Vec2f64{0, 0});   // :29
        p.colors.resize(particles_count);
        p.radius = radius;
        return p;
    }
    void mark_dirty() { dirty = true; }
    // copy device state back into the host slices (headless callers decide when)
    void sync_host() {
        if (!sim || dirty) return;
        check(ps_download(sim, &positions[0].x, &velocities[0].x), sim, "ps_download");
        check(ps_download_acc(sim, &accelerations[0].x), sim, "ps_download_acc");
    }
    void deinit() {                                               // :41-44
        if (sim) { ps_destroy(sim); sim = nullptr; }
    }
};

// src/core.zig:47-93: parameters only, scratch lives on the device.
struct SpatialHashGrid {
    size_t rows = 0, cols = 0;
    double grid_size = 0, width = 0, height = 0;
    size_t particles_count = 0;
    static SpatialHashGrid init(double width, double height, double grid_size, size_t particles_count) {
        SpatialHashGrid g;
        const int32_t wi = (int32_t)width, hi = (int32_t)height, gi = (int32_t)grid_size;   // :61-63
        if (gi <= 0) throw Error(PS_ERR_BAD_ARG, "SpatialHashGrid.init: grid_size truncates to <= 0");
        g.cols = (size_t)(wi / gi);                                                         // :66
        g.rows = (size_t)(hi / gi);                                                         // :67
        g.grid_size = grid_size; g.width = width; g.height = height; g.particles_count = particles_count;
        return g;
    }
    void deinit() {}
};

enum class ReadBack { Frame, Never };

// src/core.zig:217-268
struct Solver {
    double delta_time = 0, dt = 0, g = 0, boundary_damping = 0, particle_damping = 0, target_fps = 0;
    size_t sub_steps = 0;
    bool t_acc = false;                         // :220
    // backend options (not in the reference)
    int precision = PS_F64;
    bool collisions = true;
    int device = 0;
    ReadBack read_back = ReadBack::Frame;       // Frame: refresh particles.positions after every call (GUI); Never: headless

    static Solver init(double target_fps, size_t sub_steps, double g, double boundary_damping, double particle_damping) {
        Solver s;
        s.target_fps = target_fps;
        s.delta_time = 1 / target_fps;                              // :227
        s.dt = s.delta_time / (double)sub_steps;                    // :228
        s.sub_steps = sub_steps; s.g = g;
        s.boundary_damping = boundary_damping; s.particle_damping = particle_damping;
        return s;
    }
    void toggle_acc() { t_acc = !t_acc; }                           // :239-241

    void updatePhysics(Particles& particles, double width, double height, SpatialHashGrid& shg) {   // :245-268
        ps_config c{};
        c.struct_size = sizeof(ps_config);
        c.device = device;
        c.count = particles.positions.size();
        c.radius = particles.radius;
        c.width = width; c.height = height; c.grid_size = shg.grid_size;
        c.target_fps = target_fps; c.sub_steps = sub_steps; c.g = g;
        c.boundary_damping = boundary_damping; c.particle_damping = particle_damping;
        c.precision = precision; c.collisions = collisions ? 1 : 0;
        if (!particles.sim || std::memcmp(&c, &particles.bound, sizeof(c)) != 0) {
            if (particles.sim) { particles.sync_host(); ps_destroy(particles.sim); particles.sim = nullptr; }
            check(ps_create(&c, &particles.sim), nullptr, "ps_create");
            particles.bound = c;
            particles.dirty = true;
        }
        ps_sim* sim = particles.sim;
        if (particles.dirty) {
            check(ps_upload(sim, &particles.positions[0].x, &particles.velocities[0].x), sim, "ps_upload");
            particles.dirty = false;
        }
        check(ps_set_acc(sim, t_acc ? 1 : 0), sim, "ps_set_acc");
        check(ps_step(sim, 1), sim, "ps_step");
        if (read_back == ReadBack::Frame)
            check(ps_download(sim, &particles.positions[0].x, nullptr), sim, "ps_download");
    }
};

// src/core.zig:394-418.  `seed` replaces *std.Random.DefaultPrng (wall-clock seeded in the
// reference, src/main.zig:45): same distribution, reproducible counter-based stream.
struct Spawner {
    static void spawnRandom(Particles& par, uint64_t seed, double width, double height, double min, double max) {
        const uint64_t base = splitmix64(seed);
        const double r = par.radius;
        for (size_t i = 0; i < par.positions.size(); ++i) {
            par.positions[i].x = r + (width - r) * uniform01(base, i, 0);     // :404
            par.positions[i].y = r + (height - r) * uniform01(base, i, 1);    // :405
            par.velocities[i].x = min + (max - min) * uniform01(base, i, 2);  // :406
            par.velocities[i].y = min + (max - min) * uniform01(base, i, 3);  // :407
        }
        par.mark_dirty();
        assignRainbowColors(par);
    }
    static void assignRainbowColors(Particles& par) {                         // :411-417
        const size_t count = par.positions.size();
        for (size_t i = 0; i < count; ++i) {
            const float hue = (float)i / (float)count * 360.0f;
            par.colors[i] = hsvToSdlColor(hue, 1.0f, 1.0f, 255);
        }
    }
};

}  // namespace core_cuda
