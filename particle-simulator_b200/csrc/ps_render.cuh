// ps_render.cuh — what main.zig does with the particles after the step, as device kernels.
//
// The reference draws every particle with one SDL_RenderCopyF of a circle texture (src/main.zig:102-107 ->
// src/gui.zig:83-86): rect {x - r, y - r, 2r, 2r} in f32, tinted with particles.colors[i].  No graphics stack
// exists on this box, so the window becomes a headless buffer:
//   export_rects   the per-particle draw data exactly as the draw loop computes it (f32 casts included)
//   rainbow        Spawner.assignRainbowColors (src/core.zig:411-417) + colors.hsvToSdlColor (src/colors.zig:6-47)
//   splat          a rasteriser for the headless image: every pixel whose centre lies inside a particle's disc takes
//                  the colour of the LAST such particle in index order (the reference paints in index order, later
//                  over earlier); hard-edged discs instead of the anti-aliased circle.png sprite.
#pragma once
#include "ps_common.cuh"

namespace ps {

// src/main.zig:102-107 + src/gui.zig:83-86
template <class Real>
__global__ void __launch_bounds__(256)
export_rects_kernel(const Real* __restrict__ px, const Real* __restrict__ py, uint64_t n, float radius, float4* __restrict__ rects) {
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    const float dia = 2.0f * radius;                                    // gui.zig:84
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        const float x = (float)px[i], y = (float)py[i];                 // @floatCast, main.zig:104-105
        rects[i] = make_float4(x - radius, y - radius, dia, dia);       // .x, .y, .w, .h  gui.zig:85
    }
}

// src/colors.zig:6-47 in f32, one rounded operation at a time
__device__ __forceinline__ uchar4 hsv_to_rgba(float h, float s, float v, unsigned char a) {
    const float c = v * s;
    const float x = c * (1.0f - fabsf(fmodf(h / 60.0f, 2.0f) - 1.0f));
    const float m = v - c;
    float r = 0.f, g = 0.f, b = 0.f;
    if (h < 60.0f) { r = c; g = x; }
    else if (h < 120.0f) { r = x; g = c; }
    else if (h < 180.0f) { g = c; b = x; }
    else if (h < 240.0f) { g = x; b = c; }
    else if (h < 300.0f) { r = x; b = c; }
    else { r = c; b = x; }
    return make_uchar4((unsigned char)((r + m) * 255.0f), (unsigned char)((g + m) * 255.0f), (unsigned char)((b + m) * 255.0f), a);
}

// Spawner.assignRainbowColors, src/core.zig:411-417
__global__ void __launch_bounds__(256)
rainbow_kernel(uint64_t n, uchar4* __restrict__ colors) {
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        const float hue = __fmul_rn(__fdiv_rn((float)i, (float)n), 360.0f);
        colors[i] = hsv_to_rgba(hue, 1.0f, 1.0f, 255);
    }
}

// pass 1: owner[pixel] = max over covering particles of (index + 1)
template <class Real>
__global__ void __launch_bounds__(256)
splat_owner_kernel(const Real* __restrict__ px, const Real* __restrict__ py, uint64_t n, float radius,
                   uint32_t w, uint32_t h, uint32_t* __restrict__ owner) {
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    const float r2 = radius * radius;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        const float x = (float)px[i], y = (float)py[i];
        const int x0 = max(0, (int)floorf(x - radius)), x1 = min((int)w - 1, (int)floorf(x + radius));
        const int y0 = max(0, (int)floorf(y - radius)), y1 = min((int)h - 1, (int)floorf(y + radius));
        for (int yy = y0; yy <= y1; ++yy)
            for (int xx = x0; xx <= x1; ++xx) {
                const float dx = ((float)xx + 0.5f) - x, dy = ((float)yy + 0.5f) - y;
                if (dx * dx + dy * dy <= r2) atomicMax(owner + (size_t)yy * w + xx, (uint32_t)i + 1u);
            }
    }
}

// pass 2: colour of the owning particle, black (0,0,0,0 — main.zig's clear colour, :141-146) where nothing was drawn
__global__ void __launch_bounds__(256)
splat_resolve_kernel(const uint32_t* __restrict__ owner, const uchar4* __restrict__ colors, uint64_t pixels, uchar4* __restrict__ image) {
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t p = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; p < pixels; p += stride) {
        const uint32_t o = owner[p];
        image[p] = o ? colors[o - 1u] : make_uchar4(0, 0, 0, 0);
    }
}

}  // namespace ps
