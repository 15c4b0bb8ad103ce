/* particles_cuda.h — C ABI of libparticles_cuda, the B200 (sm_100a) replacement for the
 * per-frame particle step of J-A-G-A-N/Particle-Simulator (src/core.zig).
 *
 * The reference has no FFI of its own: its boundary is the Zig API of src/core.zig as
 * main.zig consumes it (src/main.zig:9-13,63-75,102-109,124).  Every entry point below
 * names the reference item it stands in for (file:line into /root/reference).  The Zig
 * shim that keeps core.zig's type and method names over this ABI is zig/core_cuda.zig;
 * the C++ mirror used by the headless driver is host/core_cuda.hpp (both inside the
 * package directory); INTEGRATION.md shows the build.zig change.
 *
 * Conventions
 *   - plain C types only; every function returns a ps_status (0 = ok, < 0 = error)
 *     unless stated; nothing aborts inside the library.
 *   - host arrays are interleaved (x,y) pairs, byte-identical to Zig's []@Vector(2,f64)
 *     (src/core.zig:5,14-16); they stay owned by the caller.
 *   - one handle = one owner thread.  ps_step* are stream-ordered and return before the
 *     GPU finishes; ps_download*, ps_contact_pairs, ps_get_stats and ps_sync wait.
 *   - particle identity is the array index (src/core.zig:411-417, src/main.zig:103):
 *     downloads are always in original index order.
 *   - there is no CPU fallback: without a usable CUDA device ps_create fails with
 *     PS_ERR_NO_DEVICE.
 */
#ifndef PARTICLES_CUDA_H
#define PARTICLES_CUDA_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define PS_ABI_VERSION 1

typedef struct ps_sim ps_sim; /* opaque: Particles + Solver + SpatialHashGrid state on one GPU */

typedef enum ps_status {
    PS_OK = 0,
    PS_ERR_BAD_ARG = -1,       /* null pointer, zero grid, size mismatch, ... */
    PS_ERR_OOM = -2,           /* device or pinned-host allocation failed (Zig: error.OutOfMemory, src/core.zig:23-26,76-79) */
    PS_ERR_CUDA = -3,          /* any other CUDA runtime error; text in ps_last_error */
    PS_ERR_NCCL = -4,          /* reserved for the multi-GPU slab path */
    PS_ERR_CELL_OVERFLOW = -5, /* a grid cell held > 255 particles: the reference's u8 counter would wrap or panic (src/core.zig:51,132) */
    PS_ERR_CAPACITY = -6,      /* contact list larger than the configured capacity; state is left at the failing sub-step's post-bounce point.
                                  Like PS_ERR_CELL_OVERFLOW it is sticky: every later step returns it, until ps_upload / ps_spawn
                                  gives the handle a new state (which clears it) */
    PS_ERR_NO_DEVICE = -7,     /* no CUDA device / library built without one: the product has no CPU path */
    PS_ERR_MIGRATION = -8      /* slab path: a particle crossed a whole neighbouring slab in one sub-step */
} ps_status;

typedef enum ps_precision {
    PS_F64 = 0, /* core.zig as written (Vec2f64, src/core.zig:5) */
    PS_F32 = 1  /* same operation order, every f64 replaced by f32 (north_star's fp32 SoA layout) */
} ps_precision;

/* Scene description = the arguments main.zig passes to Particles.init, Solver.init,
 * SpatialHashGrid.init and updatePhysics (src/main.zig:63-75,109).  Zero-initialise, set
 * struct_size = sizeof(ps_config), then fill. */
typedef struct ps_config {
    uint32_t struct_size;
    int32_t  device;            /* CUDA device ordinal */
    uint64_t count;             /* Particles.init particles_count       src/core.zig:22   (main.zig: 13 000) */
    double   radius;            /* Particles.init radius                src/core.zig:22   (2.0) */
    double   width, height;     /* updatePhysics width/height           src/core.zig:245  (720, 640); also SpatialHashGrid.init :60 */
    double   grid_size;         /* SpatialHashGrid.init grid_size       src/core.zig:60   (2*radius = 4.0) */
    double   target_fps;        /* Solver.init target_fps               src/core.zig:226  (60) */
    uint64_t sub_steps;         /* Solver.init sub_steps                                  (8) */
    double   g;                 /* Solver.init g                                          (50) */
    double   boundary_damping;  /* Solver.init boundary_damping                           (0.1) */
    double   particle_damping;  /* Solver.init particle_damping                           (0.9) */
    int32_t  precision;         /* ps_precision */
    int32_t  collisions;        /* 1 = faithful updatePhysics (src/core.zig:262-266 run every sub-step);
                                   0 = integrate + wall bounce only (stop after :261) — the mode
                                   BASELINE.json's integration configs name; the reference has no such switch */
    uint32_t fuse_substeps;     /* integrate-only scenes: advance up to this many sub-steps per kernel pass in
                                   registers (bit-identical; 0 or 1 = one pass per sub-step, the roofline run) */
    uint32_t contacts_per_particle; /* capacity of the contact lists' overflow pool, entries per particle (0 = default 12) */
    uint64_t slab_capacity_percent; /* ps_slab_* only: particles one rank can hold, in percent of the average share
                                       count / nranks (0 = default 125, sized for near-uniform scenes; a scene that piles up
                                       in some slabs needs more, up to 100 * nranks; scenes of at most 4 Mi particles always
                                       get room for all of them); too little is reported as PS_ERR_CAPACITY */
    uint64_t reserved[3];
} ps_config;

typedef struct ps_stats {
    uint64_t substeps_done;     /* sub-steps completed since ps_create */
    uint64_t pairs_last;        /* collision_pairs.items.len of the last sub-step (src/core.zig:211) */
    uint64_t pairs_total;       /* summed over all sub-steps */
    uint64_t rounds_last;       /* parallel rounds the ordered resolve needed in the last sub-step */
    uint64_t rounds_total;
    uint32_t max_cell_count;    /* largest cell population seen (reference limit 255) */
    uint32_t reserved;
    uint64_t kernel_launches;   /* CUDA kernels this handle has launched since ps_create (host-side count) */
} ps_stats;

/* ---- lifetime ----------------------------------------------------------------------- */

/* Particles.init (src/core.zig:22-39) + Solver.init (:226-238) + SpatialHashGrid.init
 * (:60-85) on device cfg->device.  Positions, velocities and accelerations start at zero
 * (:27-29); gravity starts off (t_acc = false, :220). */
int ps_create(const ps_config* cfg, ps_sim** out);

/* Particles.deinit (src/core.zig:41-44) + SpatialHashGrid.deinit (:87-93). NULL is ignored. */
void ps_destroy(ps_sim* sim);

/* ---- state in / out --------------------------------------------------------------------
 * Stand in for direct access to particles.positions / .velocities (src/core.zig:14-15),
 * which Spawner.spawnRandom writes (:403-408) and main.zig reads every frame (:102-107). */
int ps_upload(ps_sim* sim, const double* pos_xy, const double* vel_xy);      /* 2*count doubles each */
int ps_upload_f32(ps_sim* sim, const float* pos_xy, const float* vel_xy);
int ps_download(ps_sim* sim, double* pos_xy, double* vel_xy);                /* either may be NULL */
int ps_download_f32(ps_sim* sim, float* pos_xy, float* vel_xy);              /* pos = the GUI's @floatCast, src/main.zig:104-105 */
/* Spawner.spawnRandom (src/core.zig:395-410) on the device: x ~ U[radius, width), y ~ U[radius, height),
 * velocity components ~ U[vel_min, vel_max), from a counter-based generator (splitmix64 of seed and particle
 * index; the reference's std.Random stream is seeded from the wall clock and is not reproduced).  Bit-identical
 * to the host generator in scenes.py / core_cuda.hpp, so no host arrays are needed for large scenes. */
int ps_spawn(ps_sim* sim, uint64_t seed, double vel_min, double vel_max);
/* particles.accelerations (src/core.zig:16): a pure function of t_acc and g (:249-257). */
int ps_download_acc(ps_sim* sim, double* acc_xy);

/* ---- what main.zig does with the state after the step (no window on a headless box) ----------------------
 * The draw loop's per-particle data (src/main.zig:102-107 -> src/gui.zig:83-86): SDL_FRect {x - r, y - r, 2r, 2r} in f32
 * (positions and radius cast with @floatCast first), 4 floats per particle, index order. */
int ps_export_rects(ps_sim* sim, float* rects_xywh);
/* particles.colors (src/core.zig:17): r,g,b,a bytes per particle.  ps_rainbow_colors = Spawner.assignRainbowColors
 * (src/core.zig:411-417) with colors.hsvToSdlColor (src/colors.zig:6-47) on the device. */
int ps_set_colors(ps_sim* sim, const uint8_t* rgba);
int ps_rainbow_colors(ps_sim* sim);
int ps_get_colors(ps_sim* sim, uint8_t* rgba);
/* A headless frame in place of the SDL window: width x height RGBA8, cleared to (0,0,0,0) (src/main.zig:100,141-146), every
 * particle painted as a disc of its radius and colour in index order (later particles over earlier ones).  The reference blits
 * an anti-aliased sprite (assets/circle.png); this rasteriser paints hard-edged discs (pixel centre inside the circle). */
int ps_render(ps_sim* sim, uint32_t width, uint32_t height, uint8_t* rgba);

/* ---- stepping ----------------------------------------------------------------------- */

/* Solver.updatePhysics (src/core.zig:245-268) called `frames` times = frames*sub_steps sub-steps. */
int ps_step(ps_sim* sim, uint64_t frames);
/* Solver.updatePhysics on HOST slices, as the reference runs it (src/core.zig:245, particles.positions / .velocities in
 * host memory): upload pos_in + vel_in, `frames` x updatePhysics, write the new positions (and velocities) to pos_out
 * (vel_out); either output may be NULL, outputs may alias inputs.  Same result as ps_upload + ps_step + ps_download.
 * Integration-only scenes with pinned buffers (ps_host_alloc) are streamed in chunks so that the upload of one chunk,
 * the kernels of another and the write-back of a third overlap (PCIe is full duplex). */
int ps_frame_host(ps_sim* sim, const double* pos_in, const double* vel_in, double* pos_out, double* vel_out, uint64_t frames);
int ps_frame_host_f32(ps_sim* sim, const float* pos_in, const float* vel_in, float* pos_out, float* vel_out, uint64_t frames);
/* `substeps` iterations of the loop body at src/core.zig:247-267 (benchmarks count sub-steps). */
int ps_step_substeps(ps_sim* sim, uint64_t substeps);
/* Solver.toggle_acc (src/core.zig:239-241): flips gravity and the collision response mode. */
int ps_toggle_acc(ps_sim* sim);
int ps_set_acc(ps_sim* sim, int on);
int ps_get_acc(const ps_sim* sim);   /* returns 0/1, or < 0 on NULL */

/* Wait for all queued work; returns any deferred device-side error (cell overflow, capacity). */
int ps_sync(ps_sim* sim);
/* Queue work on the caller's cudaStream_t instead of the handle's own stream (NULL = back to own). */
int ps_set_stream(ps_sim* sim, void* cuda_stream);

/* ---- inspection --------------------------------------------------------------------- */

/* SpatialHashGrid.rows / .cols as init derives them (src/core.zig:61-67). */
int ps_grid_shape(const ps_sim* sim, uint64_t* rows, uint64_t* cols);
/* Solver.dt (src/core.zig:228) widened to double. */
double ps_dt(const ps_sim* sim);
/* Ask for shg.collision_pairs (src/core.zig:55,203) of every following sub-step to be kept. */
int ps_capture_pairs(ps_sim* sim, int on);
/* The ordered list {particle_i, particle_j} of the last sub-step, reference order
 * (src/core.zig:164-209).  Copies min(cap, count) entries (2 uint64 each); *count = full length. */
int ps_contact_pairs(ps_sim* sim, uint64_t* out_pairs, uint64_t cap, uint64_t* count);
int ps_get_stats(ps_sim* sim, ps_stats* out);
/* Raw device pointers of the structure-of-arrays state (element type per cfg.precision), in
 * original index order, valid until ps_destroy; for callers that keep data on the GPU. */
int ps_device_state(ps_sim* sim, void** px, void** py, void** vx, void** vy);

/* ---- errors / misc ------------------------------------------------------------------ */

const char* ps_last_error(const ps_sim* sim); /* sim may be NULL: last ps_create failure on this thread */
const char* ps_status_string(int status);
int ps_abi_version(void);

/* Pinned host memory for the caller's position/velocity slices, so that uploads and the
 * per-frame read-back run at PCIe speed (std.heap.c_allocator's role, src/main.zig:47). */
int ps_host_alloc(void** ptr, uint64_t bytes);
int ps_host_free(void* ptr);

/* ---- kernels on caller-owned device memory -------------------------------------------
 * The integrate loop (src/core.zig:248-260) fused with _applyBoundary (:290-318) for
 * `substeps` sub-steps on SoA device arrays of `count` elements each (16-byte aligned,
 * padded to a multiple of 4 elements), queued on `cuda_stream`.  `cfg` supplies radius,
 * width, height, fps, sub_steps, g and damping; t_acc is the gravity flag. */
int ps_integrate_bounce_f32(const ps_config* cfg, int t_acc, float* px, float* py, float* vx, float* vy,
                            uint64_t count, uint64_t substeps, void* cuda_stream);
int ps_integrate_bounce_f64(const ps_config* cfg, int t_acc, double* px, double* py, double* vx, double* vy,
                            uint64_t count, uint64_t substeps, void* cuda_stream);

/* SpatialHashGrid.computeStartArray (src/core.zig:137-142) on caller-owned device memory:
 * start_array[i] = grid_array[0] + ... + grid_array[i-1] for i in [0, cells].  start_array holds
 * cells + 1 words; the extra last word is the particle total, which the reference substitutes for
 * the last cell's end (:168).  start_array may be the same buffer as grid_array.  The per-tile
 * partial sums (one word per 4096 cells) live in scratch memory that this call allocates and
 * frees on `cuda_stream` (cudaMallocAsync / cudaFreeAsync). */
int ps_compute_start_array(const uint32_t* grid_array, uint32_t* start_array, uint64_t cells, void* cuda_stream);

/* No counterpart in the reference: ps_compute_start_array for callers that must not allocate.
 * Some users of this library cannot let a device allocation happen inside the call: a render loop
 * like main.zig's, paced at 60 frames per second, where an allocation may stall the whole GPU for an
 * unpredictable time; or a host process that handed all device memory to its own pool up front (a
 * PyTorch process, an engine with a fixed arena), where a late cudaMalloc can simply fail.  This
 * variant computes exactly what ps_compute_start_array computes and calls no allocator of any kind:
 * its partial sums go to a 64 KiB array built into the library image (every GPU that loads the
 * library holds that array until the process exits).
 *   - limit: cells <= PS_STATIC_SCAN_MAX_CELLS = 67 108 864 (16 384 tiles of 4096 cells); larger sizes return
 *     PS_ERR_CAPACITY with a message in ps_last_error(NULL) and touch nothing.
 *   - the first call on each device looks up the array's device address and remembers it; make that first call
 *     with the device current and before starting any stream capture.  The remembered address does not survive
 *     cudaDeviceReset(): a process that resets a device must not call this entry point on it afterwards.
 *   - all calls on a device share the one array.  Nothing in it has to be cleared between calls (each
 *     call overwrites what it later reads), but two calls that overlap on the GPU — different streams,
 *     or different threads — race on it and both results are undefined.  Calls on one stream are safe.
 * Inputs and outputs are the caller's, as for ps_compute_start_array; in-place use is allowed. */
#define PS_STATIC_SCAN_MAX_CELLS 67108864ull
int ps_compute_start_array_static(const uint32_t* grid_array, uint32_t* start_array, uint64_t cells, void* cuda_stream);

/* ---- one scene across the GPUs of one box (collision scenes) ------------------------
 * No counterpart in the reference (single process, single thread).  One process per GPU; rank k owns
 * the particles whose grid column (src/core.zig:104) lies in its slab of whole cell columns and
 * exchanges migrants, a one-column halo and the boundary particles' states with its x-neighbours: by direct
 * stores into the neighbour's receive buffers over NVLink where the two processes can map each other's memory
 * (CUDA IPC; the usual case on one box), by NCCL send/recv otherwise (PS_SLAB_TRANSPORT=nccl forces that; the
 * communicator is always needed for the set-up).  Results are those of ps_step on one GPU, bit for bit.  `cfg` describes the WHOLE
 * scene (count = all particles) on every rank; cfg->collisions must be 1. */
typedef struct ps_slab ps_slab;
/* ncclGetUniqueId: rank 0 calls it and hands the 128 bytes to every rank (any host-side channel). */
int ps_nccl_unique_id(void* id128);
int ps_slab_create(const ps_config* cfg, int rank, int nranks, const void* nccl_id128, ps_slab** out);
void ps_slab_destroy(ps_slab* slab);
/* Spawner.spawnRandom (src/core.zig:395-410): every rank passes the same global arrays and keeps its slab. */
int ps_slab_upload_global(ps_slab* slab, const double* pos_xy, const double* vel_xy);
/* The same distribution generated on the device by the counter-based generator of scenes.py (no host arrays). */
int ps_slab_spawn(ps_slab* slab, uint64_t seed, double vel_min, double vel_max);
int ps_slab_step(ps_slab* slab, uint64_t frames);              /* Solver.updatePhysics x frames, this rank's share */
int ps_slab_step_substeps(ps_slab* slab, uint64_t substeps);
int ps_slab_set_acc(ps_slab* slab, int on);                    /* Solver.t_acc; must be set identically on all ranks */
int ps_slab_sync(ps_slab* slab);
int ps_slab_set_stream(ps_slab* slab, void* cuda_stream);
int ps_slab_columns(const ps_slab* slab, uint64_t* x0, uint64_t* x1);   /* own grid columns [x0, x1) */
int ps_slab_owned_count(ps_slab* slab, uint64_t* count);
/* This rank's particles: original indices plus interleaved (x,y) positions and velocities. */
int ps_slab_download_owned(ps_slab* slab, uint64_t* ids, double* pos_xy, double* vel_xy, uint64_t cap, uint64_t* count);
int ps_slab_get_stats(ps_slab* slab, ps_stats* out);           /* pairs_* count the pairs whose centre this rank owns */
const char* ps_slab_last_error(const ps_slab* slab);
/* Where a sub-step's time goes on this rank (no counterpart in the reference; the measurement SURVEY 8d/8e asks for).
 * ps_slab_set_timing(slab, 1) brackets the stages of every following sub-step with CUDA events (one extra stream
 * synchronisation per sub-step, so switch it off for throughput runs); the first 8 sub-steps after switching on are
 * discarded.  ps_slab_get_timing returns microseconds per sub-step for the stages, in order: bin (integrate + count),
 * M0 exchange, sort + detect, resolve phase A + boundary check, read-back, M1 exchange, phase B, finish. */
#define PS_SLAB_TIMING_STAGES 8
int ps_slab_set_timing(ps_slab* slab, int on);
int ps_slab_get_timing(ps_slab* slab, double* stage_us, uint32_t cap, uint64_t* substeps);

/* The same slab decomposition with all ranks emulated in ONE process on ONE GPU (device copies where the
 * multi-process path has NCCL messages).  Same kernels, same protocol, same results; it exists so that the
 * multi-rank path can be tested and profiled on a single-GPU box.  No counterpart in the reference. */
typedef struct ps_slabgroup ps_slabgroup;
int ps_slabgroup_create(const ps_config* cfg, int nranks, ps_slabgroup** out);
void ps_slabgroup_destroy(ps_slabgroup* grp);
int ps_slabgroup_upload_global(ps_slabgroup* grp, const double* pos_xy, const double* vel_xy);
int ps_slabgroup_set_acc(ps_slabgroup* grp, int on);
int ps_slabgroup_step_substeps(ps_slabgroup* grp, uint64_t substeps);
int ps_slabgroup_owned_count(ps_slabgroup* grp, int rank, uint64_t* count);
int ps_slabgroup_download_owned(ps_slabgroup* grp, int rank, uint64_t* ids, double* pos_xy, double* vel_xy, uint64_t cap, uint64_t* count);
int ps_slabgroup_get_stats(ps_slabgroup* grp, int rank, ps_stats* out);
const char* ps_slabgroup_last_error(const ps_slabgroup* grp);

#ifdef __cplusplus
}
#endif
#endif /* PARTICLES_CUDA_H */
