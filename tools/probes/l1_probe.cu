// Developer probe (not part of the library): what does a load cost on B200 when the line was last touched by
//   (a) a load of this SM, (b) a store of this SM, (c) a store of this SM followed by a throw-away ld.ca,
// for plain loads, ld.cg and ld.nc?  One warp, one lane measuring with clock64, values chained so nothing overlaps.
//   nvcc -arch=sm_100a -O3 -o /tmp/l1_probe tools/probes/l1_probe.cu && /tmp/l1_probe
#include <cstdio>
#include <cuda_runtime.h>

__device__ __forceinline__ unsigned ld_plain(const unsigned* p) { unsigned v; asm volatile("ld.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory"); return v; }
__device__ __forceinline__ unsigned ld_ca(const unsigned* p) { unsigned v; asm volatile("ld.global.ca.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory"); return v; }
__device__ __forceinline__ unsigned ld_cg(const unsigned* p) { unsigned v; asm volatile("ld.global.cg.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory"); return v; }
__device__ __forceinline__ unsigned ld_nc(const unsigned* p) { unsigned v; asm volatile("ld.global.nc.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory"); return v; }
__device__ __forceinline__ void st_plain(unsigned* p, unsigned v) { asm volatile("st.global.u32 [%0], %1;" :: "l"(p), "r"(v) : "memory"); }

__device__ __forceinline__ void prefetch_l1(const void* p) { asm volatile("prefetch.global.L1 [%0];" :: "l"(p) : "memory"); }
__device__ __forceinline__ void prefetch_l2(const void* p) { asm volatile("prefetch.global.L2 [%0];" :: "l"(p) : "memory"); }

// second region of the buffer, never touched before: prefetch every line, wait, then chase
template <int WHICH>
__global__ void probe_prefetch(unsigned* buf, int n, long long* out) {
    const int lane = threadIdx.x;
    if (lane == 1) for (int i = 0; i < n; ++i) { if (WHICH == 1) prefetch_l1(buf + i * 64); else if (WHICH == 2) prefetch_l2(buf + i * 64); }
    __syncwarp();
    const long long w0 = clock64();
    while (clock64() - w0 < 20000) {}
    unsigned idx = 0;
    const long long t0 = clock64();
    for (int i = 0; i < n; ++i) idx = ld_plain(buf + idx);
    const long long t1 = clock64();
    if (lane == 0) out[0] = (t1 - t0) / n + (idx == 12345);
}

template <int MODE>   // 0 plain, 1 ca, 2 cg, 3 nc
__device__ __forceinline__ unsigned ld(const unsigned* p) { return MODE == 0 ? ld_plain(p) : MODE == 1 ? ld_ca(p) : MODE == 2 ? ld_cg(p) : ld_nc(p); }

// buf[i] holds the index of the next element to visit (a permutation with stride 64 words = 256 B: one line each)
template <int MODE>
__global__ void probe(unsigned* buf, int n, long long* out) {
    const int lane = threadIdx.x;
    unsigned idx = 0;
    long long t0, t1;
    // (0) cold: first touch
    t0 = clock64();
    for (int i = 0; i < n; ++i) idx = ld<MODE>(buf + idx);
    t1 = clock64();
    if (lane == 0) out[0] = (t1 - t0) / n;
    // (a) again: last touched by a load of this SM
    t0 = clock64();
    for (int i = 0; i < n; ++i) idx = ld<MODE>(buf + idx);
    t1 = clock64();
    if (lane == 0) out[1] = (t1 - t0) / n + (idx == 12345);
    __syncwarp();
    // (b) lane 1 rewrites every element with the same value; lane 0 then chases
    if (lane == 1) { unsigned j = 0; for (int i = 0; i < n; ++i) { const unsigned nx = ld_cg(buf + j); st_plain(buf + j, nx); j = nx; } }
    __syncwarp();
    __threadfence_block();
    t0 = clock64();
    for (int i = 0; i < n; ++i) idx = ld<MODE>(buf + idx);
    t1 = clock64();
    if (lane == 0) out[2] = (t1 - t0) / n + (idx == 12345);
    __syncwarp();
    // (c) lane 1 rewrites and touches each line with ld.ca afterwards
    if (lane == 1) { unsigned j = 0; unsigned s = 0; for (int i = 0; i < n; ++i) { const unsigned nx = ld_cg(buf + j); st_plain(buf + j, nx); s += ld_ca(buf + j); j = nx; } if (s == 7) out[7] = s; }
    __syncwarp();
    __threadfence_block();
    t0 = clock64();
    for (int i = 0; i < n; ++i) idx = ld<MODE>(buf + idx);
    t1 = clock64();
    if (lane == 0) out[3] = (t1 - t0) / n + (idx == 12345);
}

int main() {
    const int n = 256, stride = 64;
    unsigned h[n * stride];
    for (int i = 0; i < n * stride; ++i) h[i] = 0;
    for (int i = 0; i < n; ++i) h[i * stride] = ((i + 1) % n) * stride;
    unsigned* d; long long* o; long long ho[8];
    cudaMalloc(&d, sizeof(h)); cudaMalloc(&o, 64);
    const char* names[4] = {"ld (plain)", "ld.ca", "ld.cg", "ld.nc"};
    for (int mode = 0; mode < 4; ++mode) {
        cudaMemcpy(d, h, sizeof(h), cudaMemcpyHostToDevice);
        cudaMemset(o, 0, 64);
        if (mode == 0) probe<0><<<1, 32>>>(d, n, o); else if (mode == 1) probe<1><<<1, 32>>>(d, n, o);
        else if (mode == 2) probe<2><<<1, 32>>>(d, n, o); else probe<3><<<1, 32>>>(d, n, o);
        cudaMemcpy(ho, o, 64, cudaMemcpyDeviceToHost);
        printf("%-10s cycles per dependent load: cold %lld | after own-SM load %lld | after own-SM store %lld | after store + ld.ca touch %lld\n",
               names[mode], ho[0], ho[1], ho[2], ho[3]);
    }
    // a large scratch write pushes the chase buffer out of L2 first
    unsigned* big; cudaMalloc(&big, 512u << 20);
    const char* pn[3] = {"no prefetch", "prefetch.global.L1", "prefetch.global.L2"};
    for (int which = 0; which < 3; ++which) {
        for (int evict = 0; evict < 2; ++evict) {
            cudaMemcpy(d, h, sizeof(h), cudaMemcpyHostToDevice);
            if (evict) cudaMemset(big, which + 1, 512u << 20);
            cudaMemset(o, 0, 64);
            if (which == 0) probe_prefetch<0><<<1, 32>>>(d, n, o); else if (which == 1) probe_prefetch<1><<<1, 32>>>(d, n, o); else probe_prefetch<2><<<1, 32>>>(d, n, o);
            cudaMemcpy(ho, o, 64, cudaMemcpyDeviceToHost);
            printf("%-20s %s: %lld cycles per dependent plain load\n", pn[which], evict ? "(buffer evicted from L2 first)" : "(buffer in L2)", ho[0]);
        }
    }
    printf("%s\n", cudaGetErrorString(cudaDeviceSynchronize()));
    return 0;
}
