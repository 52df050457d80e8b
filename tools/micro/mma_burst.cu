// Micro-benchmark: time of one recurrence-step-sized BURST of cta_group::2 tcgen05.mma kind::f16 (M=256, N=32/64, K=16, A in
// tensor memory), from the first issue to the commit being seen, as a function of the burst length - separates the per-MMA
// cost from the fixed cost (first-MMA latency, commit -> mbarrier -> waiter).  B walks over 16 distinct chunks as in ll_lstm.
//   nvcc -gencode arch=compute_100a,code=sm_100a -std=c++20 -I autovc_b200/csrc tools/micro/mma_burst.cu -o /tmp/mma_burst
#include <cstdio>
#include <cstdint>
#include <cuda.h>
#include <cuda_runtime.h>
#include "ptx.cuh"
using namespace avc;
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); return 1; } } while (0)

template <int N, int ISSUERS>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(128, 1)
k(long long* out, int burst, int bursts, int vary_b) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* sB = smem;                 // 16 chunks x (N/2 rows x 128 B)
    __shared__ uint64_t bar;
    __shared__ uint32_t slot;
    const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0);
    const uint32_t crank = ptx::cluster_ctarank();
    for (int i = threadIdx.x; i < (16 * 8192) / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0x3c003c00u;
    if (threadIdx.x == 0) { ptx::mbar_init(&bar, ISSUERS); ptx::fence_barrier_init(); }
    if (warp == 0) ptx::tmem_alloc_pair<512>(&slot);
    ptx::fence_proxy_async_smem();
    ptx::tc_fence_before();
    __syncthreads();
    ptx::cluster_sync();
    ptx::tc_fence_after();
    const uint32_t tmem = slot;
    const uint32_t idesc = ptx::umma_idesc_f16(1, 256, N);
    constexpr int HB = N / 2 * 128;     // bytes per chunk and CTA
    if (crank == 0 && warp >= 1 && warp <= ISSUERS) {
        const int ih = warp - 1;
        const uint32_t b_sm = ptx::smem_u32(sB);
        for (int r = 0; r < bursts; ++r) {
            const long long t0 = clock64();
            for (int i = ih * (burst / ISSUERS); i < (ih + 1) * (burst / ISSUERS); ++i) {
                const uint32_t d = tmem + (uint32_t)(ih * N);
                const uint32_t chunk = vary_b ? (uint32_t)((i >> 2) & 15) : 0u;
                const uint64_t bdesc = ptx::umma_desc_sw128(b_sm + chunk * HB) + (uint64_t)((i & 3) * 2);
                ptx::umma2_f16_ts_warp(d, tmem + 128u + (uint32_t)(((i >> 2) % 12) * 32 + (i & 3) * 8), bdesc, idesc, 1u);
            }
            const long long t1 = clock64();
            ptx::umma2_commit_warp(&bar);
            ptx::mbar_wait(&bar, (uint32_t)r & 1);
            const long long t2 = clock64();
            if ((threadIdx.x & 31) == 0 && ih == 0) { out[2 * r] = t1 - t0; out[2 * r + 1] = t2 - t0; }
            const long long w0 = clock64();
            while (clock64() - w0 < 3000) {}
        }
    } else if (warp >= 1 && warp <= ISSUERS) {
        for (int r = 0; r < bursts; ++r) ptx::mbar_wait(&bar, (uint32_t)r & 1);
    }
    ptx::tc_fence_before();
    __syncthreads();
    ptx::cluster_sync();
    if (warp == 0) { ptx::tc_fence_after(); ptx::tmem_dealloc_pair<512>(tmem); }
}

int main() {
    long long* d; CK(cudaMalloc(&d, 16 * 8 * 2));
    const int smem = 16 * 8192 + 1024;
    const int bursts = 8;
    printf("cta_group::2 M=256 K=16 bf16, A in TMEM: cycles per burst (issue loop / until commit seen), median-ish of %d bursts\n", bursts);
#define RUN(N_, IS_, BURST_, VB_) do { \
        CK(cudaFuncSetAttribute(k<N_, IS_>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem)); \
        k<N_, IS_><<<2, 128, smem>>>(d, BURST_, bursts, VB_); CK(cudaDeviceSynchronize()); \
        long long h[16]; CK(cudaMemcpy(h, d, sizeof h, cudaMemcpyDeviceToHost)); \
        printf("N=%3d issuers=%d burst=%3d varyB=%d : issue %5lld %5lld %5lld  done %5lld %5lld %5lld  -> %.1f cyc/MMA marginal vs burst/2\n", N_, IS_, BURST_, VB_, \
               h[6], h[8], h[10], h[7], h[9], h[11], 0.0); } while (0)
    for (int vb = 0; vb < 2; ++vb) {
        RUN(32, 1, 16, vb); RUN(32, 1, 32, vb); RUN(32, 1, 64, vb); RUN(32, 1, 128, vb); RUN(32, 1, 256, vb);
        RUN(32, 2, 64, vb); RUN(32, 2, 128, vb);
        RUN(64, 1, 32, vb); RUN(64, 1, 64, vb); RUN(64, 1, 128, vb);
        RUN(64, 2, 64, vb); RUN(64, 2, 128, vb);
    }
    return 0;
}
