//! core_cuda.zig — drop-in replacement for src/core.zig backed by libparticles_cuda.
//!
//! Same public names and signatures as the reference's core.zig (Particles, SpatialHashGrid,
//! Solver, Spawner; init / deinit / toggle_acc / updatePhysics / spawnRandom), so main.zig and
//! gui.zig compile unchanged after `const core = @import("core_cuda.zig");` (src/main.zig:9).
//! Every method forwards to the C ABI in include/particles_cuda.h; nothing here computes.
//!
//! NOT COMPILED IN THIS REPOSITORY'S CI: the build image has no Zig toolchain (SURVEY F9).
//! Written against Zig 0.14 (the version the reference's build.zig.zon implies).  The C++ mirror
//! host/core_cuda.hpp + host/headless_main.cpp exercises the same call sequence against the same ABI.
const std = @import("std");
const sdl = @import("zsdl2");
const Color = @import("colors.zig");
const Vec2f64 = @Vector(2, f64);

pub const c = @cImport({
    @cInclude("particles_cuda.h");
});

pub const BackendError = error{ BadArgument, OutOfMemory, CudaFailure, NcclFailure, CellOverflow, ContactCapacity, NoDevice, Migration };

fn check(rc: c_int) BackendError!void {
    return switch (rc) {
        c.PS_OK => {},
        c.PS_ERR_BAD_ARG => BackendError.BadArgument,
        c.PS_ERR_OOM => BackendError.OutOfMemory,
        c.PS_ERR_NCCL => BackendError.NcclFailure,
        c.PS_ERR_CELL_OVERFLOW => BackendError.CellOverflow,
        c.PS_ERR_CAPACITY => BackendError.ContactCapacity,
        c.PS_ERR_NO_DEVICE => BackendError.NoDevice,
        c.PS_ERR_MIGRATION => BackendError.Migration,
        else => BackendError.CudaFailure,
    };
}

/// src/core.zig:13-45.  The host slices stay (the GUI draws from `positions`, src/main.zig:102-107);
/// `sim` is the backend handle, `dirty` says the host slices are newer than the device state.
pub const Particles = struct {
    positions: []Vec2f64,
    velocities: []Vec2f64,
    accelerations: []Vec2f64,
    colors: []sdl.Color,
    radius: f64,
    allocator: std.mem.Allocator,
    sim: ?*c.ps_sim = null,
    bound: c.ps_config = std.mem.zeroes(c.ps_config),
    dirty: bool = true,
    const Self = @This();

    pub fn init(allocator: std.mem.Allocator, particles_count: usize, radius: f64) !Self {
        const positions = try allocator.alloc(Vec2f64, particles_count);
        const velocities = try allocator.alloc(Vec2f64, particles_count);
        const accelerations = try allocator.alloc(Vec2f64, particles_count);
        const colors = try allocator.alloc(sdl.Color, particles_count);
        @memset(positions, @splat(0.0));
        @memset(velocities, @splat(0.0));
        @memset(accelerations, @splat(0.0));
        return Self{
            .positions = positions,
            .velocities = velocities,
            .accelerations = accelerations,
            .colors = colors,
            .radius = radius,
            .allocator = allocator,
        };
    }

    /// Call after writing positions/velocities by hand (Spawner does it for you).
    pub fn markDirty(self: *Self) void {
        self.dirty = true;
    }

    /// Copy the device state back into the host slices (headless callers decide when).
    pub fn syncHost(self: *Self) !void {
        if (self.sim == null or self.dirty) return;
        try check(c.ps_download(self.sim, @ptrCast(self.positions.ptr), @ptrCast(self.velocities.ptr)));
        try check(c.ps_download_acc(self.sim, @ptrCast(self.accelerations.ptr)));
    }

    pub fn deinit(self: *Self) void {
        if (self.sim) |s| c.ps_destroy(s);
        self.sim = null;
        self.allocator.free(self.positions);
        self.allocator.free(self.velocities);
        self.allocator.free(self.accelerations);
        self.allocator.free(self.colors);
    }
};

/// src/core.zig:47-93: parameters only; the grid's scratch arrays live on the GPU.
pub const SpatialHashGrid = struct {
    rows: usize,
    cols: usize,
    grid_size: f64,
    const Self = @This();

    pub fn init(allocator: std.mem.Allocator, width: f64, height: f64, grid_size: f64, particles_count: usize) !Self {
        _ = allocator;
        _ = particles_count;
        const width_i32: i32 = @intFromFloat(width);
        const height_i32: i32 = @intFromFloat(height);
        const grid_size_i32: i32 = @intFromFloat(grid_size);
        return Self{
            .rows = @intCast(@divFloor(height_i32, grid_size_i32)),
            .cols = @intCast(@divFloor(width_i32, grid_size_i32)),
            .grid_size = grid_size,
        };
    }

    pub fn deinit(self: *Self) void {
        _ = self;
    }
};

/// src/core.zig:217-268
pub const Solver = struct {
    delta_time: f64,
    sub_steps: usize,
    t_acc: bool = false,
    boundary_damping: f64,
    particle_damping: f64,
    g: f64,
    dt: f64,
    target_fps: f64,
    // backend options (not in the reference)
    precision: c_int = c.PS_F64,
    collisions: bool = true,
    device: c_int = 0,
    read_back_every_frame: bool = true, // false = headless: call particles.syncHost() yourself
    const Self = @This();

    pub fn init(target_fps: f64, sub_steps: usize, g: f64, boundary_damping: f64, particle_damping: f64) Self {
        const delta_time = 1 / target_fps;
        const dt = delta_time / @as(f64, @floatFromInt(sub_steps));
        return Self{
            .delta_time = delta_time,
            .sub_steps = sub_steps,
            .dt = dt,
            .g = g,
            .boundary_damping = boundary_damping,
            .particle_damping = particle_damping,
            .target_fps = target_fps,
        };
    }

    pub fn toggle_acc(self: *Self) void {
        self.t_acc = !self.t_acc;
    }

    pub fn updatePhysics(self: *Self, particles: *Particles, width: f64, height: f64, shg: *SpatialHashGrid) !void {
        var cfg = std.mem.zeroes(c.ps_config);
        cfg.struct_size = @sizeOf(c.ps_config);
        cfg.device = self.device;
        cfg.count = particles.positions.len;
        cfg.radius = particles.radius;
        cfg.width = width;
        cfg.height = height;
        cfg.grid_size = shg.grid_size;
        cfg.target_fps = self.target_fps;
        cfg.sub_steps = self.sub_steps;
        cfg.g = self.g;
        cfg.boundary_damping = self.boundary_damping;
        cfg.particle_damping = self.particle_damping;
        cfg.precision = self.precision;
        cfg.collisions = if (self.collisions) 1 else 0;

        if (particles.sim == null or !std.mem.eql(u8, std.mem.asBytes(&cfg), std.mem.asBytes(&particles.bound))) {
            if (particles.sim) |s| {
                try particles.syncHost();
                c.ps_destroy(s);
                particles.sim = null;
            }
            var handle: ?*c.ps_sim = null;
            try check(c.ps_create(&cfg, &handle));
            particles.sim = handle;
            particles.bound = cfg;
            particles.dirty = true;
        }
        if (particles.dirty) {
            try check(c.ps_upload(particles.sim, @ptrCast(particles.positions.ptr), @ptrCast(particles.velocities.ptr)));
            particles.dirty = false;
        }
        try check(c.ps_set_acc(particles.sim, if (self.t_acc) 1 else 0));
        try check(c.ps_step(particles.sim, 1));
        if (self.read_back_every_frame) {
            try check(c.ps_download(particles.sim, @ptrCast(particles.positions.ptr), null));
        }
    }
};

/// Stand-in for src/core.zig:394-418.  Spawning is host work that happens once; this version draws the
/// same distribution (x in [radius, width), y in [radius, height), velocity components in [min, max))
/// and then flags the slices so the next updatePhysics uploads them.
pub const Spawner = struct {
    fn uniform(rng: *std.Random.DefaultPrng, lo: f64, hi: f64) f64 {
        const t = rng.random().float(f64);
        return std.math.lerp(lo, hi, t);
    }

    pub fn spawnRandom(par: *Particles, rng: *std.Random.DefaultPrng, width: f64, height: f64, min: f64, max: f64) !void {
        var i: usize = 0;
        while (i < par.positions.len) : (i += 1) {
            const px = uniform(rng, par.radius, width);
            const py = uniform(rng, par.radius, height);
            const vx = uniform(rng, min, max);
            const vy = uniform(rng, min, max);
            par.positions[i] = .{ px, py };
            par.velocities[i] = .{ vx, vy };
        }
        par.markDirty();
        assignRainbowColors(par);
    }

    pub fn assignRainbowColors(par: *Particles) void {
        const n: f32 = @floatFromInt(par.positions.len);
        for (par.colors, 0..) |*slot, idx| {
            const hue = @as(f32, @floatFromInt(idx)) / n * 360.0;
            slot.* = Color.hsvToSdlColor(hue, 1.0, 1.0, 255);
        }
    }
};
