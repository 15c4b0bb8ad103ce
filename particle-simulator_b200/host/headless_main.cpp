// headless_main.cpp — main.zig without the window.
//
// Replays the reference's call sequence exactly (src/main.zig:35-76 init, :83-114 main loop,
// :116-130 key handling) against the GPU backend: Particles.init -> Spawner.spawnRandom ->
// Solver.init -> SpatialHashGrid.init -> loop { handle events; updatePhysics }.  SDL, the renderer
// and the 60 fps sleep are gone (no graphics stack on the box); key presses come from --keys.
//
//   ps_headless [--frames N] [--particles N] [--width W] [--height H] [--radius R] [--seed S]
//               [--precision f64|f32] [--no-collisions] [--keys "g@10,space@20,right@25,escape@90"]
//               [--dump state.bin] [--readback frame|never] [--device D]
//
// Prints one JSON line: frames run, sub-steps, wall time, particle-updates/s, position checksum.
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <string>
#include <vector>

#include "core_cuda.hpp"

using namespace core_cuda;

// ---- globals, as in src/main.zig:16-33
static bool quit = false;
static bool pause_ = false;
static bool next_frame = false;
static int window_width = 720;             // :25
static int window_height = 640;            // :26
static const double boundary_damping = 0.1;   // :27
static const double particle_damping = 0.9;   // :28
static const double g = 50;                   // :29
static uint64_t frame_rate = 60;              // :21

static Particles particles;
static Solver slover;                         // sic, src/main.zig:32
static SpatialHashGrid shg;

enum class Key { escape, space, right, g };

static void handle_event(Key k) {             // src/main.zig:116-130
    switch (k) {
        case Key::escape: quit = !quit; break;
        case Key::space: pause_ = !pause_; break;
        case Key::right: next_frame = true; break;
        case Key::g: slover.toggle_acc(); break;
    }
}

static uint64_t frames_stepped = 0;

static void updateAndRender(const std::multimap<uint64_t, Key>& script, uint64_t frame) {   // src/main.zig:95-114
    auto range = script.equal_range(frame);
    for (auto it = range.first; it != range.second; ++it) handle_event(it->second);         // pollEvent loop :97-99
    // draw loop (:102-107) omitted: it only reads particles.positions, kept current by ReadBack::Frame
    if (!pause_ || next_frame) {                                                            // :108
        slover.updatePhysics(particles, (double)window_width, (double)window_height, shg);  // :109
        next_frame = false;
        ++frames_stepped;
    }
}

int main(int argc, char** argv) {
    uint64_t frames = 125, count = 13000, seed = 1;   // 125 frames = 1000 sub-steps
    double radius = 2.0;
    std::string keys, dump, readback = "frame", precision = "f64";
    bool collisions = true;
    int device = 0;
    for (int i = 1; i < argc; ++i) {
        const std::string a = argv[i];
        auto val = [&](const char* name) -> const char* {
            if (i + 1 >= argc) { std::fprintf(stderr, "%s needs a value\n", name); std::exit(2); }
            return argv[++i];
        };
        if (a == "--frames") frames = std::strtoull(val("--frames"), nullptr, 10);
        else if (a == "--particles") count = std::strtoull(val("--particles"), nullptr, 10);
        else if (a == "--width") window_width = std::atoi(val("--width"));
        else if (a == "--height") window_height = std::atoi(val("--height"));
        else if (a == "--radius") radius = std::atof(val("--radius"));
        else if (a == "--seed") seed = std::strtoull(val("--seed"), nullptr, 10);
        else if (a == "--precision") precision = val("--precision");
        else if (a == "--no-collisions") collisions = false;
        else if (a == "--keys") keys = val("--keys");
        else if (a == "--dump") dump = val("--dump");
        else if (a == "--readback") readback = val("--readback");
        else if (a == "--device") device = std::atoi(val("--device"));
        else { std::fprintf(stderr, "unknown option %s\n", a.c_str()); return 2; }
    }
    std::multimap<uint64_t, Key> script;
    for (size_t p = 0; p < keys.size();) {
        size_t q = keys.find(',', p);
        if (q == std::string::npos) q = keys.size();
        const std::string tok = keys.substr(p, q - p);
        const size_t at = tok.find('@');
        if (at == std::string::npos) { std::fprintf(stderr, "bad --keys token %s\n", tok.c_str()); return 2; }
        const std::string name = tok.substr(0, at);
        const uint64_t f = std::strtoull(tok.c_str() + at + 1, nullptr, 10);
        Key k;
        if (name == "g") k = Key::g; else if (name == "space") k = Key::space;
        else if (name == "right") k = Key::right; else if (name == "escape") k = Key::escape;
        else { std::fprintf(stderr, "unknown key %s\n", name.c_str()); return 2; }
        script.emplace(f, k);
        p = q + 1;
    }

    try {
        // ---- init(), src/main.zig:63-75
        const size_t particles_count = count;                                        // :63
        const double particles_radius = radius;                                      // :64
        particles = Particles::init(particles_count, particles_radius);              // :65
        const double max = (double)(window_height < window_width ? window_height : window_width);   // :67
        const double min = -max;                                                     // :69
        Spawner::spawnRandom(particles, seed, window_width, window_height, min, max);   // :70
        slover = Solver::init((double)frame_rate, 8, g, boundary_damping, particle_damping);   // :72
        slover.precision = precision == "f32" ? PS_F32 : PS_F64;
        slover.collisions = collisions;
        slover.device = device;
        slover.read_back = readback == "never" ? ReadBack::Never : ReadBack::Frame;
        const double grid_size = 2.0 * particles_radius;                             // :74
        shg = SpatialHashGrid::init(window_width, window_height, grid_size, particles_count);   // :75

        // ---- main loop, src/main.zig:88-92 (maintainFPS sleep dropped)
        const auto t0 = std::chrono::steady_clock::now();
        for (uint64_t frame = 0; frame < frames && !quit; ++frame) updateAndRender(script, frame);
        if (particles.sim) check(ps_sync(particles.sim), particles.sim, "ps_sync");
        const auto t1 = std::chrono::steady_clock::now();
        particles.sync_host();

        const double secs = std::chrono::duration<double>(t1 - t0).count();
        double sx = 0, sy = 0;
        for (const auto& p : particles.positions) { sx += p.x; sy += p.y; }
        ps_stats st{};
        if (particles.sim) ps_get_stats(particles.sim, &st);
        if (!dump.empty()) {
            FILE* f = std::fopen(dump.c_str(), "wb");
            if (!f) { std::fprintf(stderr, "cannot open %s\n", dump.c_str()); return 1; }
            const uint64_t n = particles.positions.size();
            std::fwrite(&n, sizeof(n), 1, f);
            std::fwrite(particles.positions.data(), sizeof(Vec2f64), n, f);
            std::fwrite(particles.velocities.data(), sizeof(Vec2f64), n, f);
            std::fclose(f);
        }
        std::printf("{\"frames\": %llu, \"substeps\": %llu, \"particles\": %llu, \"seconds\": %.6f, "
                    "\"particle_updates_per_s\": %.6g, \"gravity\": %s, \"pairs_last\": %llu, \"rounds_last\": %llu, "
                    "\"sum_x\": %.17g, \"sum_y\": %.17g}\n",
                    (unsigned long long)frames_stepped, (unsigned long long)st.substeps_done, (unsigned long long)count, secs,
                    secs > 0 ? (double)count * (double)st.substeps_done / secs : 0.0, slover.t_acc ? "true" : "false",
                    (unsigned long long)st.pairs_last, (unsigned long long)st.rounds_last, sx, sy);
        particles.deinit();                                                          // deinit(), :77-81
        shg.deinit();
    } catch (const Error& e) {
        std::fprintf(stderr, "ps_headless: %s (status %d)\n", e.what(), e.status);
        return 1;
    }
    return 0;
}
