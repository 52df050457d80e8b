// Micro-benchmark: latency of 8 x 16-byte loads (one "chunk" per thread, 512 B per warp instruction) of data that
// other SMs wrote (a) in a previous kernel, (b) a few thousand cycles ago in this kernel, for several load flavours.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); return 1; } } while (0)

__device__ __forceinline__ uint4 ld_flavour(const uint4* p, int f) {
    uint4 v;
    switch (f) {
    case 0: asm volatile("ld.relaxed.gpu.global.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory"); break;
    case 1: asm volatile("ld.volatile.global.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory"); break;
    case 2: asm volatile("ld.global.cg.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory"); break;
    case 3: asm volatile("ld.global.cv.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory"); break;
    case 4: asm volatile("ld.global.ca.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory"); break;
    default: asm volatile("ld.acquire.gpu.global.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory"); break;
    }
    return v;
}

// buf: [rounds][chunks=16][8 planes][128 rows] uint4.  mode 0: read data of a previous kernel; mode 1: write then read.
__global__ void k(uint4* buf, unsigned int* bar, long long* out, int flavour, int mode, int rounds, int loaders) {
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int nctas = gridDim.x;
    for (int r = 0; r < rounds; ++r) {
        uint4* base = buf + (size_t)r * 16 * 8 * 128;
        if (mode == 1) {
            // every CTA writes its share (like the LSTM epilogue: 256 threads x 16 B), then a grid barrier, then a delay
            if (tid < 256) {
                const int plane = (blockIdx.x % 64) * 2 + (tid >> 7);
                uint4 v = make_uint4(r, tid, blockIdx.x, 1);
                asm volatile("st.relaxed.gpu.global.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(base + plane * 128 + (tid & 127)), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
            }
        }
        __syncthreads();
        if (tid == 0) {
            __threadfence();
            atomicAdd(bar, 1u);
            while (*(volatile unsigned int*)bar < (unsigned)(nctas * (r + 1))) {}
        }
        __syncthreads();
        long long t0 = clock64();
        while (clock64() - t0 < 4000) {}
        if (warp < loaders) {
            const int chunk = (warp + blockIdx.x) & 15;
            const uint4* p = base + (size_t)chunk * 8 * 128 + (warp & 3) * 32 + lane;
            uint4 v[8];
            long long a = clock64();
#pragma unroll
            for (int i = 0; i < 8; ++i) v[i] = ld_flavour(p + i * 128, flavour);
            unsigned s = 0;
#pragma unroll
            for (int i = 0; i < 8; ++i) s += v[i].x + v[i].y + v[i].z + v[i].w;
            long long b = clock64();
            if (s == 0xdeadbeef) printf("x");
            if (lane == 0 && warp == 0) out[(size_t)blockIdx.x * rounds + r] = b - a;
        }
        __syncthreads();
    }
}

int main() {
    const int rounds = 32, nctas = 128;
    uint4* buf; unsigned int* bar; long long* out;
    size_t n = (size_t)rounds * 16 * 8 * 128;
    CK(cudaMalloc(&buf, n * 16)); CK(cudaMalloc(&bar, 4)); CK(cudaMalloc(&out, sizeof(long long) * nctas * rounds));
    long long* h = new long long[nctas * rounds];
    const char* names[] = {"ld.relaxed.gpu", "ld.volatile", "ld.global.cg", "ld.global.cv", "ld.global.ca", "ld.acquire.gpu"};
    for (int mode = 0; mode < 2; ++mode)
        for (int loaders = 16; loaders >= 1; loaders /= 4)
            for (int f = 0; f < 6; ++f) {
                CK(cudaMemset(buf, 0x11, n * 16)); CK(cudaMemset(bar, 0, 4));
                CK(cudaDeviceSynchronize());
                void* args[] = {&buf, &bar, &out, &f, &mode, (void*)&rounds, &loaders};
                CK(cudaLaunchCooperativeKernel((void*)k, dim3(nctas), dim3(512), args, 0, 0));
                CK(cudaDeviceSynchronize());
                CK(cudaMemcpy(h, out, sizeof(long long) * nctas * rounds, cudaMemcpyDeviceToHost));
                long long mn = 1 << 30, mx = 0, sum = 0; int cnt = 0;
                for (int c = 0; c < nctas; ++c) for (int r = 4; r < rounds; ++r) { long long v = h[c * rounds + r]; mn = v < mn ? v : mn; mx = v > mx ? v : mx; sum += v; ++cnt; }
                printf("mode %d (%s) loaders/CTA %2d %-15s: min %lld avg %lld max %lld cycles for 8 x LDG.128\n", mode, mode ? "fresh" : "prev-kernel", loaders, names[f], mn, sum / cnt, mx);
            }
    return 0;
}
