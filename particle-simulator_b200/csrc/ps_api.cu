// ps_api.cu — the C ABI of libparticles_cuda (include/particles_cuda.h).
//
// One ps_sim = the reference's Particles + Solver + SpatialHashGrid (src/core.zig:13-45,
// 47-93, 217-238) living on one B200.  No CPU path exists in this library: every entry
// point that computes launches CUDA kernels, and ps_create fails without a device.
#include <cuda_runtime.h>

#include <cstring>
#include <exception>
#include <memory>
#include <mutex>
#include <new>
#include <string>

#include "particles_cuda.h"
#include "ps_collide.cuh"
#include "ps_common.cuh"
#include "ps_integrate.cuh"
#include "ps_render.cuh"
#include "ps_slab.cuh"
#include "ps_transfer.cuh"

using ps::Error;

static thread_local std::string g_create_error;

struct ps_sim {
    ps_config cfg{};
    Error err;
    int device = 0;
    bool t_acc = false;            // Solver.t_acc, src/core.zig:220
    bool stepped_with_acc = false; // what particles.accelerations currently holds (Q6)
    uint64_t n = 0, n_pad = 0;
    uint32_t rows = 0, cols = 0;
    cudaStream_t own_stream = nullptr, stream = nullptr;
    void* stage = nullptr;         // device staging for pageable host buffers
    uint64_t stage_bytes = 0;
    uint64_t substeps_done = 0;
    int integrate_variant = 0;
    int transfer_mode = 0;         // upload/download: 0 = kernels address pinned host buffers directly, 1 = always DMA via staging;
                                   // ps_frame_host streaming: copy engines unless 2 (direct addressing)
    uint64_t launches = 0;         // kernels launched by this handle

    virtual ~ps_sim() {}
    virtual int upload(const void* pos, const void* vel, bool host_f64) = 0;
    virtual int download(void* pos, void* vel, bool host_f64) = 0;
    virtual int step(uint64_t substeps) = 0;
    virtual int frame_host(const void* pos_in, const void* vel_in, void* pos_out, void* vel_out, bool host_f64, uint64_t frames) = 0;
    virtual int device_state(void** px, void** py, void** vx, void** vy) = 0;
    virtual int contact_pairs(uint64_t* out, uint64_t cap, uint64_t* count) = 0;
    virtual int capture_pairs(bool on) = 0;
    virtual int stats(ps_stats* out) = 0;
    virtual int check_deferred() = 0;
    virtual double dt() const = 0;
    virtual int spawn(uint64_t seed, double vmin, double vmax) = 0;
    virtual int export_rects(float* rects) = 0;
    virtual int colors_io(const uint8_t* in, uint8_t* out, bool rainbow) = 0;
    virtual int render(uint32_t w, uint32_t h, uint8_t* rgba) = 0;

    int ensure_stage(uint64_t bytes) {
        if (bytes <= stage_bytes) return PS_OK;
        if (stage) { cudaFree(stage); stage = nullptr; stage_bytes = 0; }
        PS_CUDA_TRY(err, cudaMalloc(&stage, bytes));
        stage_bytes = bytes;
        return PS_OK;
    }
    int fail(int status, const std::string& text) { err.status = status; err.text = text; return status; }
};

namespace {

struct DeviceGuard {
    int prev = -1;
    explicit DeviceGuard(int dev) { cudaGetDevice(&prev); if (prev != dev) cudaSetDevice(dev); else prev = -1; }
    ~DeviceGuard() { if (prev >= 0) cudaSetDevice(prev); }
};

template <class Real>
struct SimT final : ps_sim {
    Real *px = nullptr, *py = nullptr, *vx = nullptr, *vy = nullptr;
    ps::Collider<Real> col;

    ~SimT() override {
        DeviceGuard g(device);
        if (own_stream) cudaStreamSynchronize(own_stream);
        col.destroy();
        cudaFree(px); cudaFree(py); cudaFree(vx); cudaFree(vy);
        cudaFree(colors); cudaFree(owner); cudaFree(image);
        cudaFree(stage);
        for (int i = 0; i < kPipeStreams; ++i) { if (pipe[i]) cudaStreamDestroy(pipe[i]); if (pipe_done[i]) cudaEventDestroy(pipe_done[i]); cudaFree(pipe_stage[i]); }
        if (pipe_go) cudaEventDestroy(pipe_go);
        if (own_stream) cudaStreamDestroy(own_stream);
    }

    int init() {
        DeviceGuard g(device);
        PS_CUDA_TRY(err, cudaStreamCreateWithFlags(&own_stream, cudaStreamNonBlocking));
        stream = own_stream;
        n_pad = ps::round_up(n ? n : 1, 4);
        const size_t bytes = n_pad * sizeof(Real);
        PS_CUDA_TRY(err, cudaMalloc(&px, bytes));
        PS_CUDA_TRY(err, cudaMalloc(&py, bytes));
        PS_CUDA_TRY(err, cudaMalloc(&vx, bytes));
        PS_CUDA_TRY(err, cudaMalloc(&vy, bytes));
        // Particles.init zero-fills (src/core.zig:27-29)
        PS_CUDA_TRY(err, cudaMemsetAsync(px, 0, bytes, stream));
        PS_CUDA_TRY(err, cudaMemsetAsync(py, 0, bytes, stream));
        PS_CUDA_TRY(err, cudaMemsetAsync(vx, 0, bytes, stream));
        PS_CUDA_TRY(err, cudaMemsetAsync(vy, 0, bytes, stream));
        if (cfg.collisions) {
            const int rc = col.create(cfg, n, rows, cols, stream, err);
            if (rc != PS_OK) return rc;
        }
        PS_CUDA_TRY(err, cudaStreamSynchronize(stream));
        return PS_OK;
    }

    template <class H>
    int upload_one(const H* host_xy, Real* ax, Real* ay) {
        using HP = typename ps::Pair<H>::type;
        void* alias = nullptr;
        const HP* src;
        if (transfer_mode == 0 && ps::host_ptr_is_pinned(host_xy, &alias)) {
            src = reinterpret_cast<const HP*>(alias);
        } else {
            const uint64_t bytes = n * sizeof(HP);
            const int rc = ensure_stage(bytes);
            if (rc != PS_OK) return rc;
            PS_CUDA_TRY(err, cudaMemcpyAsync(stage, host_xy, bytes, cudaMemcpyHostToDevice, stream));
            src = reinterpret_cast<const HP*>(stage);
        }
        ps::deinterleave_kernel<H, Real><<<ps::transfer_grid(n), 256, 0, stream>>>(src, ax, ay, n);
        ++launches;
        PS_CUDA_TRY(err, cudaGetLastError());
        return PS_OK;
    }

    int upload(const void* pos, const void* vel, bool host_f64) override {
        DeviceGuard g(device);
        if (n == 0) return PS_OK;
        int rc = PS_OK;
        if (cfg.collisions) {   // a partial upload (positions only / velocities only) must find the other array current
            rc = col.materialize_soa(px, py, vx, vy, stream, err, &launches);
            if (rc != PS_OK) return rc;
        }
        if (pos) rc = host_f64 ? upload_one((const double*)pos, px, py) : upload_one((const float*)pos, px, py);
        if (rc != PS_OK) return rc;
        if (vel) rc = host_f64 ? upload_one((const double*)vel, vx, vy) : upload_one((const float*)vel, vx, vy);
        if (rc != PS_OK) return rc;
        if (cfg.collisions && (pos || vel)) { col.invalidate_records(); PS_CUDA_TRY(err, col.clear_flags(stream)); err = Error{}; }
        // the caller may reuse or free its buffers as soon as we return
        PS_CUDA_TRY(err, cudaStreamSynchronize(stream));
        return PS_OK;
    }

    template <class H>
    int download_one(H* host_xy, const Real* ax, const Real* ay) {
        using HP = typename ps::Pair<H>::type;
        void* alias = nullptr;
        ++launches;
        if (transfer_mode == 0 && ps::host_ptr_is_pinned(host_xy, &alias)) {
            ps::interleave_kernel<Real, H><<<ps::transfer_grid(n), 256, 0, stream>>>(ax, ay, reinterpret_cast<HP*>(alias), n);
            PS_CUDA_TRY(err, cudaGetLastError());
        } else {
            const uint64_t bytes = n * sizeof(HP);
            const int rc = ensure_stage(bytes);
            if (rc != PS_OK) return rc;
            ps::interleave_kernel<Real, H><<<ps::transfer_grid(n), 256, 0, stream>>>(ax, ay, reinterpret_cast<HP*>(stage), n);
            PS_CUDA_TRY(err, cudaGetLastError());
            PS_CUDA_TRY(err, cudaMemcpyAsync(host_xy, stage, bytes, cudaMemcpyDeviceToHost, stream));
            // wait before `stage` is reused by the next array
            PS_CUDA_TRY(err, cudaStreamSynchronize(stream));
        }
        return PS_OK;
    }

    int download(void* pos, void* vel, bool host_f64) override {
        DeviceGuard g(device);
        int rc = PS_OK;
        if (n) {
            if (cfg.collisions) {   // collision scenes keep the state in cell order; rebuild index order on demand (Q9)
                rc = col.materialize_soa(px, py, vx, vy, stream, err, &launches);
                if (rc != PS_OK) return rc;
            }
            if (pos) rc = host_f64 ? download_one((double*)pos, px, py) : download_one((float*)pos, px, py);
            if (rc != PS_OK) return rc;
            if (vel) rc = host_f64 ? download_one((double*)vel, vx, vy) : download_one((float*)vel, vx, vy);
            if (rc != PS_OK) return rc;
        }
        PS_CUDA_TRY(err, cudaStreamSynchronize(stream));
        return check_deferred();
    }

    // ---- host buffers in, one or more frames, host buffers out ------------------------------------------------
    // Same result as upload + step + download.  When the scene is integration-only (particles independent) and
    // the host buffers are pinned, the particles are cut into chunks that travel through a small pool of streams:
    // while chunk i is written back to the host over PCIe, chunk i+1 is being read from it (the link is full
    // duplex) and both overlap the integrate kernels of other chunks.
    static constexpr int kPipeStreams = 4;
    cudaStream_t pipe[kPipeStreams] = {};
    cudaEvent_t pipe_done[kPipeStreams] = {}, pipe_go = nullptr;
    void* pipe_stage[kPipeStreams] = {};     // transfer_mode 1: per-stream device staging for the copy engines
    size_t pipe_stage_bytes = 0;

    template <class H>
    int frame_host_pipelined(const H* pos_in, const H* vel_in, H* pos_out, H* vel_out, uint64_t frames) {
        using HP = typename ps::Pair<H>::type;
        void *a_pi = nullptr, *a_vi = nullptr, *a_po = nullptr, *a_vo = nullptr;
        if (!ps::host_ptr_is_pinned(pos_in, &a_pi) || !ps::host_ptr_is_pinned(vel_in, &a_vi)) return -100;
        if (pos_out && !ps::host_ptr_is_pinned(pos_out, &a_po)) return -100;
        if (vel_out && !ps::host_ptr_is_pinned(vel_out, &a_vo)) return -100;
        if (!pipe_go) {
            for (int i = 0; i < kPipeStreams; ++i) {
                PS_CUDA_TRY(err, cudaStreamCreateWithFlags(&pipe[i], cudaStreamNonBlocking));
                PS_CUDA_TRY(err, cudaEventCreateWithFlags(&pipe_done[i], cudaEventDisableTiming));
            }
            PS_CUDA_TRY(err, cudaEventCreateWithFlags(&pipe_go, cudaEventDisableTiming));
        }
        const ps::StepConst<Real> k = ps::make_step_const<Real>(cfg, t_acc, rows, cols);
        const uint64_t substeps = frames * cfg.sub_steps;
        const uint64_t fuse = cfg.fuse_substeps > 1 ? cfg.fuse_substeps : 1;
        const uint64_t chunk = 2ull << 20;                               // particles per chunk (multiple of 4)
        PS_CUDA_TRY(err, cudaEventRecord(pipe_go, stream));
        for (int i = 0; i < kPipeStreams; ++i) PS_CUDA_TRY(err, cudaStreamWaitEvent(pipe[i], pipe_go, 0));
        const unsigned tg = 148 * 4;                                     // modest grids so that chunks of different streams co-run
        // Copy engines + device staging (default) or, with transfer mode 2, kernels that address the pinned host buffers
        // directly.  Measured on B200 / PCIe 5: 43 ms per 64 Mi-particle frame with the copy engines (both directions
        // busy at once), 57 ms with direct addressing.
        const bool dma = transfer_mode != 2;
        if (dma && pipe_stage_bytes < 3 * chunk * sizeof(HP)) {
            for (int i = 0; i < kPipeStreams; ++i) { cudaFree(pipe_stage[i]); pipe_stage[i] = nullptr; }
            pipe_stage_bytes = 0;
            for (int i = 0; i < kPipeStreams; ++i) PS_CUDA_TRY(err, cudaMalloc(&pipe_stage[i], 3 * chunk * sizeof(HP)));
            pipe_stage_bytes = 3 * chunk * sizeof(HP);
        }
        int slot = 0;
        for (uint64_t first = 0; first < n; first += chunk, slot = (slot + 1) % kPipeStreams) {
            const uint64_t m = n - first < chunk ? n - first : chunk;
            cudaStream_t st = pipe[slot];
            const HP* src_p = reinterpret_cast<const HP*>(a_pi) + first;
            const HP* src_v = reinterpret_cast<const HP*>(a_vi) + first;
            HP* stg = static_cast<HP*>(pipe_stage[slot]);
            if (dma) {
                PS_CUDA_TRY(err, cudaMemcpyAsync(stg, pos_in + 2 * first, m * sizeof(HP), cudaMemcpyHostToDevice, st));
                PS_CUDA_TRY(err, cudaMemcpyAsync(stg + chunk, vel_in + 2 * first, m * sizeof(HP), cudaMemcpyHostToDevice, st));
                src_p = stg; src_v = stg + chunk;
            }
            ps::deinterleave_kernel<H, Real><<<tg, 256, 0, st>>>(src_p, px + first, py + first, m);
            ps::deinterleave_kernel<H, Real><<<tg, 256, 0, st>>>(src_v, vx + first, vy + first, m);
            launches += 2;
            for (uint64_t left = substeps; left;) {
                const uint64_t now = left < fuse ? left : fuse;
                PS_CUDA_TRY(err, ps::launch_integrate_bounce<Real>(px + first, py + first, vx + first, vy + first, ps::round_up(m, 4), k, (int)now,
                                                                   integrate_variant ? integrate_variant : 3, st));
                ++launches;
                left -= now;
            }
            if (!dma) {
                if (a_po) { ps::interleave_kernel<Real, H><<<tg, 256, 0, st>>>(px + first, py + first, reinterpret_cast<HP*>(a_po) + first, m); ++launches; }
                if (a_vo) { ps::interleave_kernel<Real, H><<<tg, 256, 0, st>>>(vx + first, vy + first, reinterpret_cast<HP*>(a_vo) + first, m); ++launches; }
            } else {
                if (pos_out) {
                    ps::interleave_kernel<Real, H><<<tg, 256, 0, st>>>(px + first, py + first, stg + 2 * chunk, m); ++launches;
                    PS_CUDA_TRY(err, cudaMemcpyAsync(pos_out + 2 * first, stg + 2 * chunk, m * sizeof(HP), cudaMemcpyDeviceToHost, st));
                }
                if (vel_out) {                                           // reuses the velocity-in staging area (already consumed on this stream)
                    ps::interleave_kernel<Real, H><<<tg, 256, 0, st>>>(vx + first, vy + first, stg + chunk, m); ++launches;
                    PS_CUDA_TRY(err, cudaMemcpyAsync(vel_out + 2 * first, stg + chunk, m * sizeof(HP), cudaMemcpyDeviceToHost, st));
                }
            }
        }
        PS_CUDA_TRY(err, cudaGetLastError());
        for (int i = 0; i < kPipeStreams; ++i) {
            PS_CUDA_TRY(err, cudaEventRecord(pipe_done[i], pipe[i]));
            PS_CUDA_TRY(err, cudaStreamWaitEvent(stream, pipe_done[i], 0));
        }
        if (substeps) stepped_with_acc = t_acc;
        substeps_done += substeps;
        PS_CUDA_TRY(err, cudaStreamSynchronize(stream));                 // the caller owns its buffers again on return
        return PS_OK;
    }

    int frame_host(const void* pos_in, const void* vel_in, void* pos_out, void* vel_out, bool host_f64, uint64_t frames) override {
        DeviceGuard g(device);
        if (!cfg.collisions && pos_in && vel_in && n >= (4ull << 20)) {
            const int rc = host_f64 ? frame_host_pipelined((const double*)pos_in, (const double*)vel_in, (double*)pos_out, (double*)vel_out, frames)
                                    : frame_host_pipelined((const float*)pos_in, (const float*)vel_in, (float*)pos_out, (float*)vel_out, frames);
            if (rc != -100) return rc;                                   // -100: buffers not pinned, take the plain route
        }
        int rc = upload(pos_in, vel_in, host_f64);
        if (rc != PS_OK) return rc;
        rc = step(frames * cfg.sub_steps);
        if (rc != PS_OK) return rc;
        return download(pos_out, vel_out, host_f64);
    }

    int step(uint64_t substeps) override {
        DeviceGuard g(device);
        if (err.status != PS_OK && err.status != PS_ERR_BAD_ARG) return err.status;   // sticky device-side failure
        const ps::StepConst<Real> k = ps::make_step_const<Real>(cfg, t_acc, rows, cols);
        if (substeps) stepped_with_acc = t_acc;
        if (!cfg.collisions) {
            const uint64_t fuse = cfg.fuse_substeps > 1 ? cfg.fuse_substeps : 1;
            uint64_t left = substeps;
            while (left) {
                const uint64_t now = left < fuse ? left : fuse;
                PS_CUDA_TRY(err, ps::launch_integrate_bounce<Real>(px, py, vx, vy, n_pad, k, (int)now, integrate_variant, stream));
                ++launches;
                left -= now;
            }
        } else {
            for (uint64_t s = 0; s < substeps; ++s) {
                const int rc = col.substep(px, py, vx, vy, k, stream, err, &launches);
                if (rc != PS_OK) return rc;
            }
        }
        substeps_done += substeps;
        return PS_OK;
    }

    int device_state(void** ox, void** oy, void** ovx, void** ovy) override {
        DeviceGuard g(device);
        if (cfg.collisions && n) {
            const int rc = col.materialize_soa(px, py, vx, vy, stream, err, &launches);
            if (rc != PS_OK) return rc;
        }
        if (ox) *ox = px; if (oy) *oy = py; if (ovx) *ovx = vx; if (ovy) *ovy = vy;
        return PS_OK;
    }

    int capture_pairs(bool on) override {
        if (!cfg.collisions) return fail(PS_ERR_BAD_ARG, "ps_capture_pairs: handle was created with collisions = 0");
        col.capture = on;
        return PS_OK;
    }

    int contact_pairs(uint64_t* out, uint64_t cap, uint64_t* count) override {
        DeviceGuard g(device);
        if (!cfg.collisions) return fail(PS_ERR_BAD_ARG, "ps_contact_pairs: handle was created with collisions = 0");
        PS_CUDA_TRY(err, cudaStreamSynchronize(stream));
        const int rc = check_deferred();
        if (rc != PS_OK) return rc;
        return col.read_pairs(out, cap, count, stream, err);
    }

    int stats(ps_stats* out) override {
        DeviceGuard g(device);
        PS_CUDA_TRY(err, cudaStreamSynchronize(stream));
        std::memset(out, 0, sizeof(*out));
        out->substeps_done = substeps_done;
        out->kernel_launches = launches;
        if (cfg.collisions) {
            const int rc = col.read_stats(out, stream, err);
            if (rc != PS_OK) return rc;
        }
        return check_deferred();
    }

    int check_deferred() override {
        if (!cfg.collisions) return PS_OK;
        return col.check_flags(stream, err);
    }

    double dt() const override { return (double)ps::make_step_const<Real>(cfg, false, rows, cols).dt; }

    // ---- after the step: draw data and a headless frame (ps_render.cuh) -------------------------------------
    uchar4* colors = nullptr;
    uint32_t* owner = nullptr; uchar4* image = nullptr; uint64_t image_pixels = 0;

    int ensure_colors() {
        if (colors) return PS_OK;
        PS_CUDA_TRY(err, cudaMalloc((void**)&colors, (n ? n : 1) * sizeof(uchar4)));
        PS_CUDA_TRY(err, cudaMemsetAsync(colors, 0xff, (n ? n : 1) * sizeof(uchar4), stream));   // white until the caller sets them
        return PS_OK;
    }
    int current_soa() {
        if (cfg.collisions && n) return col.materialize_soa(px, py, vx, vy, stream, err, &launches);
        return PS_OK;
    }

    int export_rects(float* rects) override {
        DeviceGuard g(device);
        if (n == 0) return PS_OK;
        int rc = current_soa(); if (rc != PS_OK) return rc;
        rc = ensure_stage(n * sizeof(float4)); if (rc != PS_OK) return rc;
        ps::export_rects_kernel<Real><<<ps::transfer_grid(n), 256, 0, stream>>>(px, py, n, (float)cfg.radius, static_cast<float4*>(stage));
        ++launches;
        PS_CUDA_TRY(err, cudaMemcpyAsync(rects, stage, n * sizeof(float4), cudaMemcpyDeviceToHost, stream));
        PS_CUDA_TRY(err, cudaStreamSynchronize(stream));
        return PS_OK;
    }

    int colors_io(const uint8_t* in, uint8_t* out, bool rainbow) override {
        DeviceGuard g(device);
        int rc = ensure_colors(); if (rc != PS_OK) return rc;
        if (n == 0) return PS_OK;
        if (in) PS_CUDA_TRY(err, cudaMemcpyAsync(colors, in, n * sizeof(uchar4), cudaMemcpyHostToDevice, stream));
        if (rainbow) { ps::rainbow_kernel<<<ps::transfer_grid(n), 256, 0, stream>>>(n, colors); ++launches; }
        if (out) PS_CUDA_TRY(err, cudaMemcpyAsync(out, colors, n * sizeof(uchar4), cudaMemcpyDeviceToHost, stream));
        PS_CUDA_TRY(err, cudaStreamSynchronize(stream));
        return PS_OK;
    }

    int render(uint32_t w, uint32_t h, uint8_t* rgba) override {
        DeviceGuard g(device);
        const uint64_t pixels = (uint64_t)w * h;
        if (!pixels || !rgba) return fail(PS_ERR_BAD_ARG, "ps_render: empty image");
        if (n >= 0xffffffffull) return fail(PS_ERR_BAD_ARG, "ps_render: too many particles");
        int rc = ensure_colors(); if (rc != PS_OK) return rc;
        rc = current_soa(); if (rc != PS_OK) return rc;
        if (pixels != image_pixels) {
            cudaFree(owner); cudaFree(image); owner = nullptr; image = nullptr; image_pixels = 0;
            PS_CUDA_TRY(err, cudaMalloc((void**)&owner, pixels * sizeof(uint32_t)));
            PS_CUDA_TRY(err, cudaMalloc((void**)&image, pixels * sizeof(uchar4)));
            image_pixels = pixels;
        }
        PS_CUDA_TRY(err, cudaMemsetAsync(owner, 0, pixels * sizeof(uint32_t), stream));
        if (n) { ps::splat_owner_kernel<Real><<<ps::transfer_grid(n), 256, 0, stream>>>(px, py, n, (float)cfg.radius, w, h, owner); ++launches; }
        ps::splat_resolve_kernel<<<ps::transfer_grid(pixels), 256, 0, stream>>>(owner, colors, pixels, image);
        ++launches;
        PS_CUDA_TRY(err, cudaMemcpyAsync(rgba, image, pixels * sizeof(uchar4), cudaMemcpyDeviceToHost, stream));
        PS_CUDA_TRY(err, cudaStreamSynchronize(stream));
        return PS_OK;
    }

    int spawn(uint64_t seed, double vmin, double vmax) override {
        DeviceGuard g(device);
        if (n == 0) return PS_OK;
        uint64_t z = seed + 0x9E3779B97F4A7C15ull;           // splitmix64(seed): the generator's stream base (scenes.py uniform01)
        z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull; z = (z ^ (z >> 27)) * 0x94D049BB133111EBull; z ^= z >> 31;
        ps::spawn_soa_kernel<Real><<<148 * 16, 256, 0, stream>>>(n, z, cfg.radius, cfg.width, cfg.height, vmin, vmax, px, py, vx, vy);
        ++launches;
        PS_CUDA_TRY(err, cudaGetLastError());
        if (cfg.collisions) { col.invalidate_records(); PS_CUDA_TRY(err, col.clear_flags(stream)); err = Error{}; }
        PS_CUDA_TRY(err, cudaStreamSynchronize(stream));
        return PS_OK;
    }
};

int validate(const ps_config* cfg, uint32_t* rows, uint32_t* cols, std::string* why) {
    if (!cfg) { *why = "cfg is NULL"; return PS_ERR_BAD_ARG; }
    if (cfg->struct_size != sizeof(ps_config)) { *why = "cfg->struct_size != sizeof(ps_config) (ABI mismatch)"; return PS_ERR_BAD_ARG; }
    if (cfg->precision != PS_F64 && cfg->precision != PS_F32) { *why = "unknown precision"; return PS_ERR_BAD_ARG; }
    if (!(cfg->target_fps > 0) || cfg->sub_steps == 0) { *why = "target_fps and sub_steps must be positive"; return PS_ERR_BAD_ARG; }
    if (!(cfg->radius > 0) || !(cfg->width > 0) || !(cfg->height > 0)) { *why = "radius, width, height must be positive"; return PS_ERR_BAD_ARG; }
    if (cfg->count >= (1ull << 28) + 1 && cfg->collisions) { *why = "collision scenes hold at most 2^28 particles per GPU"; return PS_ERR_BAD_ARG; }
    *rows = *cols = 0;
    if (cfg->collisions) {
        // SpatialHashGrid.init: integer division of the truncated values (src/core.zig:61-67)
        const int32_t wi = (int32_t)cfg->width, hi = (int32_t)cfg->height, gi = (int32_t)cfg->grid_size;
        if (gi <= 0) { *why = "grid_size truncates to <= 0 (src/core.zig:63)"; return PS_ERR_BAD_ARG; }
        const int64_t c = wi / gi, r = hi / gi;
        if (c <= 0 || r <= 0) { *why = "grid has no cells"; return PS_ERR_BAD_ARG; }
        if (c * r >= (1ll << 31)) { *why = "grid has more than 2^31 cells"; return PS_ERR_BAD_ARG; }
        *cols = (uint32_t)c; *rows = (uint32_t)r;
    }
    return PS_OK;
}

template <class Real>
int create_t(const ps_config* cfg, uint32_t rows, uint32_t cols, ps_sim** out) {
    auto* s = new (std::nothrow) SimT<Real>();
    if (!s) { g_create_error = "host allocation failed"; return PS_ERR_OOM; }
    s->cfg = *cfg; s->device = cfg->device; s->n = cfg->count; s->rows = rows; s->cols = cols;
    const int rc = s->init();
    if (rc != PS_OK) { g_create_error = s->err.text; delete s; return rc; }
    *out = s;
    return PS_OK;
}

}  // namespace

extern "C" {

int ps_abi_version(void) { return PS_ABI_VERSION; }

const char* ps_status_string(int status) {
    switch (status) {
        case PS_OK: return "ok";
        case PS_ERR_BAD_ARG: return "bad argument";
        case PS_ERR_OOM: return "out of memory";
        case PS_ERR_CUDA: return "CUDA error";
        case PS_ERR_NCCL: return "NCCL error";
        case PS_ERR_CELL_OVERFLOW: return "grid cell holds more than 255 particles";
        case PS_ERR_CAPACITY: return "contact list capacity exceeded";
        case PS_ERR_NO_DEVICE: return "no CUDA device";
        case PS_ERR_MIGRATION: return "a particle crossed a whole slab in one sub-step";
        default: return "unknown status";
    }
}

const char* ps_last_error(const ps_sim* sim) { return sim ? sim->err.text.c_str() : g_create_error.c_str(); }

int ps_create(const ps_config* cfg, ps_sim** out) {
    if (!out) { g_create_error = "out is NULL"; return PS_ERR_BAD_ARG; }
    *out = nullptr;
    uint32_t rows = 0, cols = 0;
    std::string why;
    int rc = validate(cfg, &rows, &cols, &why);
    if (rc != PS_OK) { g_create_error = "ps_create: " + why; return rc; }
    int ndev = 0;
    const cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0) {
        cudaGetLastError();
        g_create_error = std::string("ps_create: no CUDA device (") + cudaGetErrorString(e) + "); this library has no CPU path";
        return PS_ERR_NO_DEVICE;
    }
    if (cfg->device < 0 || cfg->device >= ndev) { g_create_error = "ps_create: device ordinal out of range"; return PS_ERR_BAD_ARG; }
    try {   // no C++ exception may cross the C boundary
        return cfg->precision == PS_F64 ? create_t<double>(cfg, rows, cols, out) : create_t<float>(cfg, rows, cols, out);
    } catch (const std::exception& ex) { g_create_error = std::string("ps_create: ") + ex.what(); return PS_ERR_OOM; }
}

void ps_destroy(ps_sim* sim) { delete sim; }

int ps_upload(ps_sim* s, const double* pos, const double* vel) { return s ? s->upload(pos, vel, true) : PS_ERR_BAD_ARG; }
int ps_upload_f32(ps_sim* s, const float* pos, const float* vel) { return s ? s->upload(pos, vel, false) : PS_ERR_BAD_ARG; }
int ps_download(ps_sim* s, double* pos, double* vel) { return s ? s->download(pos, vel, true) : PS_ERR_BAD_ARG; }
int ps_download_f32(ps_sim* s, float* pos, float* vel) { return s ? s->download(pos, vel, false) : PS_ERR_BAD_ARG; }

int ps_spawn(ps_sim* s, uint64_t seed, double vel_min, double vel_max) { return s ? s->spawn(seed, vel_min, vel_max) : PS_ERR_BAD_ARG; }

int ps_export_rects(ps_sim* s, float* rects_xywh) { return (s && (rects_xywh || !s->n)) ? s->export_rects(rects_xywh) : PS_ERR_BAD_ARG; }
int ps_set_colors(ps_sim* s, const uint8_t* rgba) { return (s && rgba) ? s->colors_io(rgba, nullptr, false) : PS_ERR_BAD_ARG; }
int ps_rainbow_colors(ps_sim* s) { return s ? s->colors_io(nullptr, nullptr, true) : PS_ERR_BAD_ARG; }
int ps_get_colors(ps_sim* s, uint8_t* rgba) { return (s && rgba) ? s->colors_io(nullptr, rgba, false) : PS_ERR_BAD_ARG; }
int ps_render(ps_sim* s, uint32_t width, uint32_t height, uint8_t* rgba) { return s ? s->render(width, height, rgba) : PS_ERR_BAD_ARG; }

int ps_download_acc(ps_sim* s, double* acc_xy) {
    if (!s || !acc_xy) return PS_ERR_BAD_ARG;
    // accelerations = (zeroNearZero(0), zeroNearZero(g)) after a gravity step, else (0,0)  src/core.zig:249-257
    double ay = 0.0;
    if (s->stepped_with_acc) {
        ay = s->cfg.precision == PS_F32 ? (double)(float)s->cfg.g : s->cfg.g;
        if ((ay < 0 ? -ay : ay) < (s->cfg.precision == PS_F32 ? (double)1e-6f : 1e-6)) ay = 0.0;
    }
    for (uint64_t i = 0; i < s->n; ++i) { acc_xy[2 * i] = 0.0; acc_xy[2 * i + 1] = ay; }
    return PS_OK;
}

int ps_step(ps_sim* s, uint64_t frames) { return s ? s->step(frames * s->cfg.sub_steps) : PS_ERR_BAD_ARG; }
int ps_frame_host(ps_sim* s, const double* pos_in, const double* vel_in, double* pos_out, double* vel_out, uint64_t frames) {
    return s ? s->frame_host(pos_in, vel_in, pos_out, vel_out, true, frames) : PS_ERR_BAD_ARG;
}
int ps_frame_host_f32(ps_sim* s, const float* pos_in, const float* vel_in, float* pos_out, float* vel_out, uint64_t frames) {
    return s ? s->frame_host(pos_in, vel_in, pos_out, vel_out, false, frames) : PS_ERR_BAD_ARG;
}
int ps_step_substeps(ps_sim* s, uint64_t substeps) { return s ? s->step(substeps) : PS_ERR_BAD_ARG; }
int ps_toggle_acc(ps_sim* s) { if (!s) return PS_ERR_BAD_ARG; s->t_acc = !s->t_acc; return PS_OK; }
int ps_set_acc(ps_sim* s, int on) { if (!s) return PS_ERR_BAD_ARG; s->t_acc = on != 0; return PS_OK; }
int ps_get_acc(const ps_sim* s) { return s ? (s->t_acc ? 1 : 0) : PS_ERR_BAD_ARG; }

int ps_sync(ps_sim* s) {
    if (!s) return PS_ERR_BAD_ARG;
    DeviceGuard g(s->device);
    PS_CUDA_TRY(s->err, cudaStreamSynchronize(s->stream));
    return s->check_deferred();
}

int ps_set_stream(ps_sim* s, void* cuda_stream) {
    if (!s) return PS_ERR_BAD_ARG;
    DeviceGuard g(s->device);
    PS_CUDA_TRY(s->err, cudaStreamSynchronize(s->stream));
    s->stream = cuda_stream ? (cudaStream_t)cuda_stream : s->own_stream;
    return PS_OK;
}

int ps_grid_shape(const ps_sim* s, uint64_t* rows, uint64_t* cols) {
    if (!s || !rows || !cols) return PS_ERR_BAD_ARG;
    if (s->cfg.collisions) { *rows = s->rows; *cols = s->cols; return PS_OK; }
    const int32_t wi = (int32_t)s->cfg.width, hi = (int32_t)s->cfg.height, gi = (int32_t)s->cfg.grid_size;
    if (gi <= 0) return PS_ERR_BAD_ARG;
    *cols = (uint64_t)(wi / gi); *rows = (uint64_t)(hi / gi);
    return PS_OK;
}

double ps_dt(const ps_sim* s) { return s ? s->dt() : 0.0; }
int ps_capture_pairs(ps_sim* s, int on) { return s ? s->capture_pairs(on != 0) : PS_ERR_BAD_ARG; }
int ps_contact_pairs(ps_sim* s, uint64_t* out, uint64_t cap, uint64_t* count) {
    if (!s || !count || (cap && !out)) return PS_ERR_BAD_ARG;
    return s->contact_pairs(out, cap, count);
}
int ps_get_stats(ps_sim* s, ps_stats* out) { return (s && out) ? s->stats(out) : PS_ERR_BAD_ARG; }
int ps_device_state(ps_sim* s, void** px, void** py, void** vx, void** vy) { return s ? s->device_state(px, py, vx, vy) : PS_ERR_BAD_ARG; }

int ps_host_alloc(void** ptr, uint64_t bytes) {
    if (!ptr) return PS_ERR_BAD_ARG;
    const cudaError_t e = cudaHostAlloc(ptr, bytes, cudaHostAllocPortable | cudaHostAllocMapped);
    if (e != cudaSuccess) { cudaGetLastError(); *ptr = nullptr; return e == cudaErrorMemoryAllocation ? PS_ERR_OOM : PS_ERR_NO_DEVICE; }
    return PS_OK;
}
int ps_host_free(void* ptr) { return cudaFreeHost(ptr) == cudaSuccess ? PS_OK : PS_ERR_CUDA; }

// Tuning hook (not part of the stable ABI): pick the integrate kernel variant, 0 = auto.
int ps_debug_set_integrate_variant(ps_sim* s, int variant) { if (!s) return PS_ERR_BAD_ARG; s->integrate_variant = variant; return PS_OK; }
int ps_debug_set_transfer_mode(ps_sim* s, int mode) { if (!s) return PS_ERR_BAD_ARG; s->transfer_mode = mode; return PS_OK; }

static int integrate_raw_check(const ps_config* cfg, const void* a, const void* b, const void* c, const void* d, uint64_t count) {
    if (!cfg || cfg->struct_size != sizeof(ps_config)) return PS_ERR_BAD_ARG;
    if (count && (!a || !b || !c || !d)) return PS_ERR_BAD_ARG;
    if (((uintptr_t)a | (uintptr_t)b | (uintptr_t)c | (uintptr_t)d) & 15) return PS_ERR_BAD_ARG;
    if (!(cfg->target_fps > 0) || cfg->sub_steps == 0) return PS_ERR_BAD_ARG;
    return PS_OK;
}

int ps_integrate_bounce_f32(const ps_config* cfg, int t_acc, float* px, float* py, float* vx, float* vy,
                            uint64_t count, uint64_t substeps, void* cuda_stream) {
    const int rc = integrate_raw_check(cfg, px, py, vx, vy, count);
    if (rc != PS_OK) return rc;
    const ps::StepConst<float> k = ps::make_step_const<float>(*cfg, t_acc != 0, 0, 0);
    const uint64_t fuse = cfg->fuse_substeps > 1 ? cfg->fuse_substeps : 1;
    for (uint64_t left = substeps; left;) {
        const uint64_t now = left < fuse ? left : fuse;
        if (ps::launch_integrate_bounce<float>(px, py, vx, vy, ps::round_up(count, 4), k, (int)now, 0, (cudaStream_t)cuda_stream) != cudaSuccess) return PS_ERR_CUDA;
        left -= now;
    }
    return PS_OK;
}

int ps_integrate_bounce_f64(const ps_config* cfg, int t_acc, double* px, double* py, double* vx, double* vy,
                            uint64_t count, uint64_t substeps, void* cuda_stream) {
    const int rc = integrate_raw_check(cfg, px, py, vx, vy, count);
    if (rc != PS_OK) return rc;
    const ps::StepConst<double> k = ps::make_step_const<double>(*cfg, t_acc != 0, 0, 0);
    const uint64_t fuse = cfg->fuse_substeps > 1 ? cfg->fuse_substeps : 1;
    for (uint64_t left = substeps; left;) {
        const uint64_t now = left < fuse ? left : fuse;
        if (ps::launch_integrate_bounce<double>(px, py, vx, vy, ps::round_up(count, 4), k, (int)now, 0, (cudaStream_t)cuda_stream) != cudaSuccess) return PS_ERR_CUDA;
        left -= now;
    }
    return PS_OK;
}

int ps_compute_start_array(const uint32_t* grid_array, uint32_t* start_array, uint64_t cells, void* cuda_stream) {
    if (!start_array || (cells && !grid_array)) return PS_ERR_BAD_ARG;
    cudaStream_t st = (cudaStream_t)cuda_stream;
    uint32_t* sums = nullptr;
    const cudaError_t e = cudaMallocAsync((void**)&sums, (ps::scan_tiles(cells) + 1) * sizeof(uint32_t), st);
    if (e != cudaSuccess) { cudaGetLastError(); return e == cudaErrorMemoryAllocation ? PS_ERR_OOM : PS_ERR_CUDA; }
    const cudaError_t k = ps::exclusive_scan_u32(grid_array, start_array, cells, sums, start_array + cells, nullptr, st);
    const cudaError_t f = cudaFreeAsync(sums, st);
    return (k == cudaSuccess && f == cudaSuccess) ? PS_OK : PS_ERR_CUDA;
}

// ---- slab-decomposed collision scenes (one process per GPU) ---------------------------------------------
}  // extern "C"

struct ps_slab {
    virtual ~ps_slab() {}
    virtual int upload_global(const double* pos, const double* vel) = 0;
    virtual int spawn(uint64_t seed, double vmin, double vmax) = 0;
    virtual int step(uint64_t substeps) = 0;
    virtual int set_acc(bool on) = 0;
    virtual int sync() = 0;
    virtual int set_stream(void* st) = 0;
    virtual int columns(uint64_t* x0, uint64_t* x1) const = 0;
    virtual int owned_count(uint64_t* n) = 0;
    virtual int download_owned(uint64_t* ids, double* pos, double* vel, uint64_t cap, uint64_t* count) = 0;
    virtual int stats(ps_stats* out) = 0;
    virtual const char* last_error() const = 0;
    virtual uint64_t sub_steps() const = 0;
    virtual int set_timing(bool on) = 0;
    virtual int get_timing(double* us, uint32_t cap, uint64_t* substeps) = 0;
};

namespace {
template <class Real>
struct SlabT final : ps_slab {
    ps::SlabCollider<Real> c;
    ~SlabT() override { DeviceGuard g(c.device); c.destroy(); }
    int upload_global(const double* pos, const double* vel) override { DeviceGuard g(c.device); return c.upload_global(pos, vel); }
    int spawn(uint64_t seed, double vmin, double vmax) override { DeviceGuard g(c.device); return c.spawn(seed, vmin, vmax); }
    int step(uint64_t substeps) override {
        DeviceGuard g(c.device);
        for (uint64_t s = 0; s < substeps; ++s) { const int rc = c.substep(); if (rc != PS_OK) return rc; }
        return PS_OK;
    }
    int set_acc(bool on) override { c.t_acc = on; return PS_OK; }
    int sync() override { DeviceGuard g(c.device); return c.sync(); }
    int set_stream(void* st) override {
        DeviceGuard g(c.device);
        cudaStreamSynchronize(c.stream);
        c.stream = st ? (cudaStream_t)st : c.own_stream;
        return PS_OK;
    }
    int columns(uint64_t* x0, uint64_t* x1) const override { *x0 = c.g.x0; *x1 = c.g.x1; return PS_OK; }
    int owned_count(uint64_t* n) override { DeviceGuard g(c.device); cudaStreamSynchronize(c.stream); *n = c.last.own_e - c.last.own_b; return PS_OK; }
    int download_owned(uint64_t* ids, double* pos, double* vel, uint64_t cap, uint64_t* count) override {
        DeviceGuard g(c.device);
        return c.download_owned(ids, pos, vel, cap, count);
    }
    int stats(ps_stats* out) override { DeviceGuard g(c.device); return c.stats(out); }
    const char* last_error() const override { return c.err.text.c_str(); }
    uint64_t sub_steps() const override { return c.cfg.sub_steps; }
    int set_timing(bool on) override { DeviceGuard g(c.device); c.set_timing(on); return PS_OK; }
    int get_timing(double* us, uint32_t cap, uint64_t* substeps) override {
        DeviceGuard g(c.device);
        cudaStreamSynchronize(c.stream);
        const uint64_t nsub = c.timed_substeps();
        if (substeps) *substeps = nsub;
        for (uint32_t i = 0; i < cap; ++i) us[i] = (i + 1 < (uint32_t)c.kTicks && nsub) ? c.acc_ms[i + 1] * 1e3 / (double)nsub : 0.0;
        return PS_OK;
    }
};

template <class Real>
int slab_create_t(const ps_config* cfg, int rank, int nranks, const void* id, uint32_t rows, uint32_t cols, ps_slab** out) {
    auto* s = new (std::nothrow) SlabT<Real>();
    if (!s) { g_create_error = "host allocation failed"; return PS_ERR_OOM; }
    DeviceGuard g(cfg->device);
    const int rc = s->c.create(*cfg, rank, nranks, static_cast<const ps::NcclUniqueId*>(id), rows, cols);
    if (rc != PS_OK) { g_create_error = s->c.err.text; delete s; return rc; }
    *out = s;
    return PS_OK;
}
}  // namespace

extern "C" {

int ps_nccl_unique_id(void* id128) {
    if (!id128) return PS_ERR_BAD_ARG;
    const ps::NcclApi& api = ps::nccl_api();
    if (!api.ok) { g_create_error = api.error; return PS_ERR_NCCL; }
    return api.GetUniqueId(static_cast<ps::NcclUniqueId*>(id128)) == ps::kNcclSuccess ? PS_OK : PS_ERR_NCCL;
}

int ps_slab_create(const ps_config* cfg, int rank, int nranks, const void* nccl_id128, ps_slab** out) {
    if (!out) { g_create_error = "out is NULL"; return PS_ERR_BAD_ARG; }
    *out = nullptr;
    uint32_t rows = 0, cols = 0;
    std::string why;
    int rc = validate(cfg, &rows, &cols, &why);
    if (rc != PS_OK) { g_create_error = "ps_slab_create: " + why; return rc; }
    if (!cfg->collisions) { g_create_error = "ps_slab_create: slabs are for collision scenes (integration-only scenes shard by index, no exchange)"; return PS_ERR_BAD_ARG; }
    if (nranks < 1 || rank < 0 || rank >= nranks) { g_create_error = "ps_slab_create: bad rank / nranks"; return PS_ERR_BAD_ARG; }
    if (cfg->count >= (1ull << 32)) { g_create_error = "ps_slab_create: particle indices are 32-bit"; return PS_ERR_BAD_ARG; }
    int ndev = 0;
    const cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0) { cudaGetLastError(); g_create_error = "ps_slab_create: no CUDA device; this library has no CPU path"; return PS_ERR_NO_DEVICE; }
    if (cfg->device < 0 || cfg->device >= ndev) { g_create_error = "ps_slab_create: device ordinal out of range"; return PS_ERR_BAD_ARG; }
    try {
        return cfg->precision == PS_F64 ? slab_create_t<double>(cfg, rank, nranks, nccl_id128, rows, cols, out)
                                        : slab_create_t<float>(cfg, rank, nranks, nccl_id128, rows, cols, out);
    } catch (const std::exception& ex) { g_create_error = std::string("ps_slab_create: ") + ex.what(); return PS_ERR_OOM; }
}
void ps_slab_destroy(ps_slab* s) { delete s; }
int ps_slab_upload_global(ps_slab* s, const double* pos, const double* vel) { return (s && pos && vel) ? s->upload_global(pos, vel) : PS_ERR_BAD_ARG; }
int ps_slab_spawn(ps_slab* s, uint64_t seed, double vmin, double vmax) { return s ? s->spawn(seed, vmin, vmax) : PS_ERR_BAD_ARG; }
int ps_slab_step(ps_slab* s, uint64_t frames) { return s ? s->step(frames * s->sub_steps()) : PS_ERR_BAD_ARG; }
int ps_slab_step_substeps(ps_slab* s, uint64_t substeps) { return s ? s->step(substeps) : PS_ERR_BAD_ARG; }
int ps_slab_set_acc(ps_slab* s, int on) { return s ? s->set_acc(on != 0) : PS_ERR_BAD_ARG; }
int ps_slab_sync(ps_slab* s) { return s ? s->sync() : PS_ERR_BAD_ARG; }
int ps_slab_set_stream(ps_slab* s, void* st) { return s ? s->set_stream(st) : PS_ERR_BAD_ARG; }
int ps_slab_columns(const ps_slab* s, uint64_t* x0, uint64_t* x1) { return (s && x0 && x1) ? s->columns(x0, x1) : PS_ERR_BAD_ARG; }
int ps_slab_owned_count(ps_slab* s, uint64_t* n) { return (s && n) ? s->owned_count(n) : PS_ERR_BAD_ARG; }
int ps_slab_download_owned(ps_slab* s, uint64_t* ids, double* pos, double* vel, uint64_t cap, uint64_t* count) {
    return (s && count) ? s->download_owned(ids, pos, vel, cap, count) : PS_ERR_BAD_ARG;
}
int ps_slab_get_stats(ps_slab* s, ps_stats* out) { return (s && out) ? s->stats(out) : PS_ERR_BAD_ARG; }
const char* ps_slab_last_error(const ps_slab* s) { return s ? s->last_error() : g_create_error.c_str(); }
int ps_slab_set_timing(ps_slab* s, int on) { return s ? s->set_timing(on != 0) : PS_ERR_BAD_ARG; }
int ps_slab_get_timing(ps_slab* s, double* stage_us, uint32_t cap, uint64_t* substeps) {
    return (s && (stage_us || !cap)) ? s->get_timing(stage_us, cap, substeps) : PS_ERR_BAD_ARG;
}

// ---- several slab ranks emulated in one process ---------------------------------------------------------------
}  // extern "C"

struct ps_slabgroup {
    virtual ~ps_slabgroup() {}
    virtual int upload_global(const double* pos, const double* vel) = 0;
    virtual int set_acc(bool on) = 0;
    virtual int step(uint64_t substeps) = 0;
    virtual int owned_count(int rank, uint64_t* n) = 0;
    virtual int download_owned(int rank, uint64_t* ids, double* pos, double* vel, uint64_t cap, uint64_t* count) = 0;
    virtual int stats(int rank, ps_stats* out) = 0;
    virtual const char* last_error() const = 0;
};

namespace {
template <class Real>
struct SlabGroupT final : ps_slabgroup {
    ps::SlabGroupLocal<Real> grp;
    int device = 0;
    std::string text;
    int note(int rc, const ps::Error& e) { if (rc != PS_OK) text = e.text; return rc; }
    bool bad(int rank) const { return rank < 0 || rank >= (int)grp.ranks.size(); }
    int upload_global(const double* pos, const double* vel) override {
        DeviceGuard g(device);
        for (auto* r : grp.ranks) { const int rc = r->upload_global(pos, vel); if (rc != PS_OK) return note(rc, r->err); }
        return PS_OK;
    }
    int set_acc(bool on) override { for (auto* r : grp.ranks) r->t_acc = on; return PS_OK; }
    int step(uint64_t substeps) override {
        DeviceGuard g(device);
        for (uint64_t s = 0; s < substeps; ++s) { const int rc = grp.substep(); if (rc != PS_OK) return note(rc, grp.err); }
        for (auto* r : grp.ranks) { const int rc = r->sync(); if (rc != PS_OK) return note(rc, r->err); }
        return PS_OK;
    }
    int owned_count(int rank, uint64_t* n) override { if (bad(rank)) return PS_ERR_BAD_ARG; *n = grp.ranks[rank]->last.own_e - grp.ranks[rank]->last.own_b; return PS_OK; }
    int download_owned(int rank, uint64_t* ids, double* pos, double* vel, uint64_t cap, uint64_t* count) override {
        if (bad(rank)) return PS_ERR_BAD_ARG;
        DeviceGuard g(device);
        return note(grp.ranks[rank]->download_owned(ids, pos, vel, cap, count), grp.ranks[rank]->err);
    }
    int stats(int rank, ps_stats* out) override {
        if (bad(rank)) return PS_ERR_BAD_ARG;
        DeviceGuard g(device);
        return note(grp.ranks[rank]->stats(out), grp.ranks[rank]->err);
    }
    const char* last_error() const override { return text.c_str(); }
};

template <class Real>
int slabgroup_create_t(const ps_config* cfg, int nranks, uint32_t rows, uint32_t cols, ps_slabgroup** out) {
    auto* s = new (std::nothrow) SlabGroupT<Real>();
    if (!s) { g_create_error = "host allocation failed"; return PS_ERR_OOM; }
    s->device = cfg->device;
    DeviceGuard g(cfg->device);
    const int rc = s->grp.create(*cfg, nranks, rows, cols);
    if (rc != PS_OK) { g_create_error = s->grp.err.text; delete s; return rc; }
    *out = s;
    return PS_OK;
}
}  // namespace

extern "C" {

int ps_slabgroup_create(const ps_config* cfg, int nranks, ps_slabgroup** out) {
    if (!out) { g_create_error = "out is NULL"; return PS_ERR_BAD_ARG; }
    *out = nullptr;
    uint32_t rows = 0, cols = 0;
    std::string why;
    int rc = validate(cfg, &rows, &cols, &why);
    if (rc != PS_OK) { g_create_error = "ps_slabgroup_create: " + why; return rc; }
    if (!cfg->collisions || nranks < 1 || cfg->count >= (1ull << 32)) { g_create_error = "ps_slabgroup_create: bad arguments"; return PS_ERR_BAD_ARG; }
    int ndev = 0;
    const cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0) { cudaGetLastError(); g_create_error = "ps_slabgroup_create: no CUDA device; this library has no CPU path"; return PS_ERR_NO_DEVICE; }
    if (cfg->device < 0 || cfg->device >= ndev) { g_create_error = "ps_slabgroup_create: device ordinal out of range"; return PS_ERR_BAD_ARG; }
    try {
        return cfg->precision == PS_F64 ? slabgroup_create_t<double>(cfg, nranks, rows, cols, out) : slabgroup_create_t<float>(cfg, nranks, rows, cols, out);
    } catch (const std::exception& ex) { g_create_error = std::string("ps_slabgroup_create: ") + ex.what(); return PS_ERR_OOM; }
}
void ps_slabgroup_destroy(ps_slabgroup* g) { delete g; }
int ps_slabgroup_upload_global(ps_slabgroup* g, const double* pos, const double* vel) { return (g && pos && vel) ? g->upload_global(pos, vel) : PS_ERR_BAD_ARG; }
int ps_slabgroup_set_acc(ps_slabgroup* g, int on) { return g ? g->set_acc(on != 0) : PS_ERR_BAD_ARG; }
int ps_slabgroup_step_substeps(ps_slabgroup* g, uint64_t substeps) { return g ? g->step(substeps) : PS_ERR_BAD_ARG; }
int ps_slabgroup_owned_count(ps_slabgroup* g, int rank, uint64_t* n) { return (g && n) ? g->owned_count(rank, n) : PS_ERR_BAD_ARG; }
int ps_slabgroup_download_owned(ps_slabgroup* g, int rank, uint64_t* ids, double* pos, double* vel, uint64_t cap, uint64_t* count) {
    return (g && count) ? g->download_owned(rank, ids, pos, vel, cap, count) : PS_ERR_BAD_ARG;
}
int ps_slabgroup_get_stats(ps_slabgroup* g, int rank, ps_stats* out) { return (g && out) ? g->stats(rank, out) : PS_ERR_BAD_ARG; }
const char* ps_slabgroup_last_error(const ps_slabgroup* g) { return g ? g->last_error() : g_create_error.c_str(); }

// ---- ps_compute_start_array_static: the same prefix sum with no allocation anywhere ------------------
//
// Scratch for the per-tile partial sums lives in the file-scope device array below instead of memory
// obtained at call time.  64 KiB per GPU that loads this library, held until the process exits.
// One word per 4096-cell tile => at most PS_STATIC_SCAN_MAX_CELLS cells per call.
#define PS_STATIC_SCAN_MAX_TILES 16384u
static_assert(PS_STATIC_SCAN_MAX_CELLS == (uint64_t)PS_STATIC_SCAN_MAX_TILES * (uint64_t)ps::SCAN_TILE, "header and scratch array disagree");
static __device__ __align__(256) uint32_t g_static_scan_sums[PS_STATIC_SCAN_MAX_TILES];

namespace {
constexpr int kMaxDevices = 64;
std::mutex g_static_scan_mutex;
uint32_t* g_static_scan_ptr[kMaxDevices] = {};   // device address of g_static_scan_sums, per device

// In host code `g_static_scan_sums` is only a handle for the runtime, not an address a kernel can
// use, and every device has its own copy of the array.  Resolve the real address once per device,
// with that device current, and remember it.
int static_scan_scratch(uint32_t** out) {
    int dev = -1;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= kMaxDevices) { cudaGetLastError(); return PS_ERR_NO_DEVICE; }
    std::lock_guard<std::mutex> lock(g_static_scan_mutex);
    if (!g_static_scan_ptr[dev]) {
        void* p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&p, g_static_scan_sums);
        if (e != cudaSuccess || !p) { cudaGetLastError(); return PS_ERR_CUDA; }
        g_static_scan_ptr[dev] = static_cast<uint32_t*>(p);
    }
    *out = g_static_scan_ptr[dev];
    return PS_OK;
}
}  // namespace

int ps_compute_start_array_static(const uint32_t* grid_array, uint32_t* start_array, uint64_t cells, void* cuda_stream) {
    if (!start_array || (cells && !grid_array)) return PS_ERR_BAD_ARG;
    if (cells > PS_STATIC_SCAN_MAX_CELLS) {
        g_create_error = "ps_compute_start_array_static: " + std::to_string(cells) + " cells exceed the built-in scratch (" +
                         std::to_string(PS_STATIC_SCAN_MAX_CELLS) + " cells); use ps_compute_start_array";
        return PS_ERR_CAPACITY;
    }
    uint32_t* sums = nullptr;
    const int rc = static_scan_scratch(&sums);      // first call on this device resolves the symbol address
    if (rc != PS_OK) return rc;
    // Every partial sum the scan reads is written earlier in the same call, so nothing left over from
    // the previous call needs clearing.
    const cudaError_t k = ps::exclusive_scan_u32(grid_array, start_array, cells, sums, start_array + cells, nullptr,
                                                 (cudaStream_t)cuda_stream);
    return k == cudaSuccess ? PS_OK : PS_ERR_CUDA;
}

}  // extern "C"
