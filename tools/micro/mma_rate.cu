// Micro-benchmark: issue rate of cta_group::2 tcgen05.mma kind::f16 (M=256) as a function of N, the A operand's home
// (tensor memory vs shared memory) and the number of independent accumulators the MMAs rotate over.
// Operands are uninitialised: only timing matters.   nvcc -gencode arch=compute_100a,code=sm_100a -I autovc_b200/csrc
#include <cstdio>
#include <cstdint>
#include <cuda.h>
#include <cuda_runtime.h>
#include "ptx.cuh"
using namespace avc;
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); return 1; } } while (0)

template <int N, int chains, bool use_ts>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(128, 1)
k(long long* out, int total, int rotate_a) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* sA = smem;                 // 8 x 16 KiB: A chunks (128 rows x 64 bf16, SW128)
    uint8_t* sB = smem + 8 * 16384;     // 16 KiB
    __shared__ uint64_t bar;
    __shared__ uint32_t slot;
    const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0);
    const uint32_t crank = ptx::cluster_ctarank();
    for (int i = threadIdx.x; i < (9 * 16384) / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0x3c003c00u;
    if (threadIdx.x == 0) { ptx::mbar_init(&bar, 1); ptx::fence_barrier_init(); }
    if (warp == 0) ptx::tmem_alloc_pair<512>(&slot);
    ptx::fence_proxy_async_smem();
    ptx::tc_fence_before();
    __syncthreads();
    ptx::cluster_sync();
    ptx::tc_fence_after();
    const uint32_t tmem = slot;
    const uint32_t idesc = ptx::umma_idesc_f16(1, 256, N);
    constexpr int dcols = chains * N;   // accumulators at columns [0, dcols); A (tensor memory) after them
    if (crank == 0 && warp == 1) {
        const long long t0 = clock64();
        const uint32_t a_sm = ptx::smem_u32(sA), b_sm = ptx::smem_u32(sB);
        const uint32_t arot = rotate_a ? 1u : 0u;
        for (int o = 0; o < total; o += 32) {
#pragma unroll
            for (int i = 0; i < 32; ++i) {
                const uint32_t d = tmem + (uint32_t)((i % chains) * N);
                const uint32_t ac = arot * (uint32_t)((i >> 2) & 7);
                const uint64_t bdesc = ptx::umma_desc_sw128(b_sm) + (uint64_t)((i & 3) * 2);
                if (use_ts) ptx::umma2_f16_ts_warp(d, tmem + (uint32_t)dcols + ((ac * 32u) & 31u) * 0u + ac * ((512 - dcols - 32) / 8 / 8 * 8) + (uint32_t)((i & 3) * 8), bdesc, idesc, 1u);
                else ptx::umma2_f16_warp(d, ptx::umma_desc_sw128(a_sm + ac * 16384u) + (uint64_t)((i & 3) * 2), bdesc, idesc, 1u);
            }
        }
        const long long t1 = clock64();
        ptx::umma2_commit_warp(&bar);
        ptx::mbar_wait(&bar, 0);
        const long long t2 = clock64();
        if ((threadIdx.x & 31) == 0) { out[0] = t1 - t0; out[1] = t2 - t0; }
    } else if (warp == 1) {
        ptx::mbar_wait(&bar, 0);
    }
    ptx::tc_fence_before();
    __syncthreads();
    ptx::cluster_sync();
    if (warp == 0) { ptx::tc_fence_after(); ptx::tmem_dealloc_pair<512>(tmem); }
}

int main() {
    long long* d; CK(cudaMalloc(&d, 16));
    const int smem = 9 * 16384 + 1024;
    const int total = 256;
    printf("cta_group::2 M=256 K=16 bf16, %d MMAs: cycles per MMA (issue loop / until commit seen)\n", total);
#define RUN(N_, CH_, TS_) do { \
        CK(cudaFuncSetAttribute(k<N_, CH_, TS_>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem)); \
        for (int rot = 0; rot < 2; ++rot) { \
            for (int rep = 0; rep < 2; ++rep) { k<N_, CH_, TS_><<<2, 128, smem>>>(d, total, rot); CK(cudaDeviceSynchronize()); } \
            long long h[2]; CK(cudaMemcpy(h, d, 16, cudaMemcpyDeviceToHost)); \
            printf("A in %s  N=%3d accumulators=%d rotateA=%d : %6.1f / %6.1f\n", TS_ ? "TMEM" : "SMEM", N_, CH_, rot, \
                   (double)h[0] / total, (double)h[1] / total); } } while (0)
    RUN(32, 1, true); RUN(32, 2, true); RUN(32, 4, true);
    RUN(64, 1, true); RUN(64, 2, true); RUN(64, 4, true);
    RUN(128, 1, true); RUN(128, 2, true); RUN(256, 1, true);
    RUN(32, 1, false); RUN(32, 2, false); RUN(32, 4, false);
    RUN(64, 1, false); RUN(64, 2, false); RUN(64, 4, false);
    RUN(128, 1, false); RUN(128, 2, false); RUN(256, 1, false);
    return 0;
}
