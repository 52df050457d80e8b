// Persistent weight-stationary LSTM recurrence on the tensor cores (sm_100a), used for the decoder's
// lstm1 (H=512) and lstm2 (H=1024, two layers run one after the other)  [reference decoder.py:37,53,65,74].
//
// Decomposition: grid = (H/16 unit slices) x (batch tiles of 128 utterances), all CTAs co-resident (cooperative
// launch, one CTA per SM).  CTA (n, bt) keeps the 64 gate rows (16 hidden units x i,f,g,o, unit-major packing) of
// W_hh resident in shared memory for the whole sequence (H=1024: 128 KiB) and its 128x16 cell state in registers.
// Per time step:
//   producer warp : waits until all H/16 slices of this batch tile have published h_{t-1} (one counter per step in
//                   global memory, release/acquire), then streams h_{t-1} [128 x H] bf16 through a shared-memory
//                   ring with TMA (same 128B-swizzled K-major tiles as the GEMM kernel)
//   MMA warp      : D[128 x 64] (TMEM, fp32) = h_{t-1} . W_slice^T with tcgen05.mma, K = H
//   epilogue warps: tcgen05.ld the gate pre-activations, add the hoisted input projection Gx_t (prefetched),
//                   sigmoid/tanh, update c in registers, store h_t (bf16) and bump the step counter.
// The input projection x_t W_ih^T + b_ih + b_hh for all t is hoisted into one dense GEMM (tc_gemm.cu) beforehand.
//
// Every spin/wait is bounded: a protocol bug traps (CUDA error) instead of hanging the GPU.
#include <cstdio>
#include <vector>
#include <cuda.h>

#include "common.cuh"
#include "launch.h"
#include "ptx.cuh"

namespace avc {

namespace {

constexpr int UNITS = 16;                 // hidden units per CTA
constexpr int NG = 4 * UNITS;             // gate rows per CTA = UMMA N
constexpr int TILE_B = 128;               // utterances per CTA = UMMA M
constexpr int KBLK = 64;                  // K elements per smem block (128 bytes)
constexpr int W_BLK_BYTES = NG * KBLK * 2;        // 8 KiB
constexpr int H_BLK_BYTES = TILE_B * KBLK * 2;    // 16 KiB
constexpr int MAX_STAGES = 8;
constexpr uint32_t TMEM_COLS = 64;

struct TcLstmArgs {
    int B, T, H;
    int Tg, gx_div;              // Gx row for frame t is t / gx_div (lstm1 reads the de-duplicated projection)
    int G_n;                     // unit slices = H / 16
    int num_kb;                  // H / 64
    int stages;
    const float* Gx;             // [B][Tg][4H] fp32, unit-major gate order, bias included
    __nv_bfloat16* Y;            // [B][T][H] bf16 output sequence (also the recurrent input)
    unsigned int* ready;         // [batch tiles][T] step counters, zeroed before launch
    uint32_t idesc;
};

__device__ __forceinline__ unsigned int ld_acquire(const unsigned int* p) {
    unsigned int v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void red_release_add(unsigned int* p, unsigned int v) {
    asm volatile("red.release.gpu.global.add.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ float fast_sigmoid(float x) { return __fdividef(1.0f, 1.0f + __expf(-x)); }
__device__ __forceinline__ float fast_tanh(float x) { return 1.0f - __fdividef(2.0f, 1.0f + __expf(2.0f * x)); }

__global__ void __launch_bounds__(192, 1)
tc_lstm_kernel(const __grid_constant__ CUtensorMap tmW, const __grid_constant__ CUtensorMap tmH, const TcLstmArgs a) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* sW = smem;                                        // num_kb x 8 KiB, resident
    uint8_t* sH = smem + a.num_kb * W_BLK_BYTES;               // stages x 16 KiB ring
    uint64_t* bars = reinterpret_cast<uint64_t*>(sH + a.stages * H_BLK_BYTES);
    uint64_t* full = bars;
    uint64_t* empty = bars + MAX_STAGES;
    uint64_t* wfull = bars + 2 * MAX_STAGES;
    uint64_t* acc_full = wfull + 1;
    uint64_t* acc_empty = wfull + 2;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(wfull + 3);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int ns = blockIdx.x;            // unit slice
    const int bt = blockIdx.y;            // batch tile
    const int b0 = bt * TILE_B;
    unsigned int* ready = a.ready + (size_t)bt * a.T;

    if (warp == 0 && lane == 0) {
        ptx::prefetch_tmap(&tmW);
        ptx::prefetch_tmap(&tmH);
        for (int i = 0; i < a.stages; ++i) { ptx::mbar_init(&full[i], 1); ptx::mbar_init(&empty[i], 1); }
        ptx::mbar_init(wfull, 1);
        ptx::mbar_init(acc_full, 1);
        ptx::mbar_init(acc_empty, 128);
        ptx::fence_barrier_init();
    }
    if (warp == 1) ptx::tmem_alloc<TMEM_COLS>(tmem_slot);
    ptx::tc_fence_before();
    __syncthreads();
    ptx::tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        // ===================== producer: resident W slice once, then h_{t-1} every step =====================
        if (lane == 0) {
            ptx::mbar_expect_tx(wfull, (uint32_t)(a.num_kb * W_BLK_BYTES));
            for (int kb = 0; kb < a.num_kb; ++kb)
                ptx::tma_load_2d(sW + kb * W_BLK_BYTES, &tmW, wfull, kb * KBLK, ns * NG);
            int stage = 0; uint32_t phase = 0;
            for (int t = 1; t < a.T; ++t) {
                // all unit slices of this batch tile must have published h_{t-1}
                bool ok = false;
                for (uint32_t spin = 0; spin < (1u << 26); ++spin) {
                    if (ld_acquire(ready + (t - 1)) >= (unsigned int)a.G_n) { ok = true; break; }
                }
                if (!ok) {
                    printf("autovc_b200: lstm step counter timed out (slice %d tile %d t %d: %u/%d)\n", ns, bt, t,
                           ld_acquire(ready + (t - 1)), a.G_n);
                    __trap();
                }
                ptx::fence_proxy_async_all();   // generic-proxy writes of h (other CTAs) -> async-proxy (TMA) reads
                for (int kb = 0; kb < a.num_kb; ++kb) {
                    ptx::mbar_wait(&empty[stage], phase ^ 1);
                    ptx::mbar_expect_tx(&full[stage], H_BLK_BYTES);
                    ptx::tma_load_3d(sH + stage * H_BLK_BYTES, &tmH, &full[stage], kb * KBLK, t - 1, b0);
                    if (++stage == a.stages) { stage = 0; phase ^= 1; }
                }
            }
        }
    } else if (warp == 1) {
        // ===================== MMA issuer =====================
        if (lane == 0) {
            ptx::mbar_wait(wfull, 0);
            int stage = 0; uint32_t phase = 0;
            for (int t = 1; t < a.T; ++t) {
                const uint32_t i = (uint32_t)(t - 1);
                ptx::mbar_wait(acc_empty, (i & 1) ^ 1);      // epilogue of step t-1 has drained the accumulator
                ptx::tc_fence_after();
                for (int kb = 0; kb < a.num_kb; ++kb) {
                    ptx::mbar_wait(&full[stage], phase);
                    ptx::tc_fence_after();
                    const uint64_t adesc = ptx::umma_desc_sw128(ptx::smem_u32(sH + stage * H_BLK_BYTES));
                    const uint64_t bdesc = ptx::umma_desc_sw128(ptx::smem_u32(sW + kb * W_BLK_BYTES));
#pragma unroll
                    for (int k = 0; k < KBLK / 16; ++k)
                        ptx::umma_f16(tmem_base, adesc + (uint64_t)(k * 2), bdesc + (uint64_t)(k * 2), a.idesc, (kb | k) != 0 ? 1u : 0u);
                    ptx::umma_commit(&empty[stage]);
                    if (++stage == a.stages) { stage = 0; phase ^= 1; }
                }
                ptx::umma_commit(acc_full);
            }
        }
    } else {
        // ===================== epilogue: gates, cell state, h_t =====================
        const int q = warp & 3;
        const int r = q * 32 + lane;
        const int b = b0 + r;
        const bool valid = b < a.B;
        const int et = threadIdx.x - 64;           // 0..127 among the epilogue threads
        const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16);
        float c[UNITS];
#pragma unroll
        for (int u = 0; u < UNITS; ++u) c[u] = 0.0f;
        const float* gx_row = a.Gx + ((size_t)(valid ? b : 0) * a.Tg) * 4 * a.H + (size_t)ns * NG;
        __nv_bfloat16* y_row = a.Y + ((size_t)(valid ? b : 0) * a.T) * a.H + (size_t)ns * UNITS;

        for (int t = 0; t < a.T; ++t) {
            // prefetch the hoisted input projection for this step (independent of the recurrence)
            float4 gx[UNITS];
            const float4* gp = reinterpret_cast<const float4*>(gx_row + (size_t)(t / a.gx_div) * 4 * a.H);
#pragma unroll
            for (int u = 0; u < UNITS; ++u) gx[u] = valid ? __ldg(gp + u) : make_float4(0.f, 0.f, 0.f, 0.f);

            uint32_t acc[4][16];
            if (t >= 1) {
                ptx::mbar_wait(acc_full, (uint32_t)(t - 1) & 1);
                ptx::tc_fence_after();
#pragma unroll
                for (int j = 0; j < 4; ++j) ptx::tmem_ld16(taddr + (uint32_t)(j * 16), acc[j]);
                ptx::tmem_ld_wait();
                ptx::tc_fence_before();
                ptx::mbar_arrive(acc_empty);
            } else {
#pragma unroll
                for (int j = 0; j < 4; ++j)
#pragma unroll
                    for (int k = 0; k < 16; ++k) acc[j][k] = 0u;
            }
            uint32_t hp[UNITS / 2];
#pragma unroll
            for (int u = 0; u < UNITS; u += 2) {
                float hv[2];
#pragma unroll
                for (int w = 0; w < 2; ++w) {
                    const int uu = u + w;
                    const int j = uu >> 2, k = (uu & 3) * 4;       // column = uu*4 + gate
                    const float ig = fast_sigmoid(__uint_as_float(acc[j][k + 0]) + gx[uu].x);
                    const float fg = fast_sigmoid(__uint_as_float(acc[j][k + 1]) + gx[uu].y);
                    const float gg = fast_tanh(__uint_as_float(acc[j][k + 2]) + gx[uu].z);
                    const float og = fast_sigmoid(__uint_as_float(acc[j][k + 3]) + gx[uu].w);
                    c[uu] = fg * c[uu] + ig * gg;
                    hv[w] = og * fast_tanh(c[uu]);
                }
                __nv_bfloat162 h2 = __floats2bfloat162_rn(hv[0], hv[1]);
                hp[u / 2] = *reinterpret_cast<uint32_t*>(&h2);
            }
            if (valid) {
                uint4* o = reinterpret_cast<uint4*>(y_row + (size_t)t * a.H);
                o[0] = make_uint4(hp[0], hp[1], hp[2], hp[3]);
                o[1] = make_uint4(hp[4], hp[5], hp[6], hp[7]);
            }
            // publish: order the generic-proxy stores before other CTAs' TMA reads, then release the step counter
            ptx::fence_proxy_async_all();
            asm volatile("bar.sync 1, 128;" ::: "memory");
            if (et == 0) {
                __threadfence();
                red_release_add(ready + t, 1u);
            }
        }
    }

    ptx::tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        ptx::tc_fence_after();
        ptx::tmem_dealloc<TMEM_COLS>(tmem_base);
    }
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
EncodeTiledFn g_encode_l = nullptr;
int g_max_smem = 0;

}  // namespace

struct TcLstmPlan {
    CUtensorMap tmW;
    std::vector<CUtensorMap> tmH;     // one per batch chunk
    std::vector<TcLstmArgs> args;
    std::vector<dim3> grids;
    int smem_bytes;
};

const char* tc_lstm_init() {
    if (g_encode_l) return nullptr;
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult qres;
    cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres);
    if (e != cudaSuccess || qres != cudaDriverEntryPointSuccess || !fn) return "cuTensorMapEncodeTiled not available from the driver";
    g_encode_l = reinterpret_cast<EncodeTiledFn>(fn);
    g_max_smem = 227 * 1024;
    if (cudaFuncSetAttribute(tc_lstm_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, g_max_smem) != cudaSuccess)
        return "cudaFuncSetAttribute(MaxDynamicSharedMemorySize) failed for tc_lstm_kernel";
    return nullptr;
}

TcLstmPlan* tc_lstm_plan_create(const void* Whh_bf16, const float* Gx, int Tg, int gx_div, void* Y, unsigned int* ready,
                                int B, int T, int H, int num_sms, const char** err) {
    static thread_local char msg[256];
    auto bad = [&](const char* m) -> TcLstmPlan* { if (err) *err = m; return nullptr; };
    if (!g_encode_l) return bad("tc_lstm_init() not called");
    if (H % 64 || H < 64) return bad("hidden size must be a multiple of 64");
    const int G_n = H / UNITS, num_kb = H / KBLK;
    if (G_n > num_sms) return bad("hidden size too large for one co-resident grid");
    const int w_bytes = num_kb * W_BLK_BYTES;
    int stages = (g_max_smem - 1024 - 512 - w_bytes) / H_BLK_BYTES;
    if (stages > MAX_STAGES) stages = MAX_STAGES;
    if (stages > num_kb) stages = num_kb;
    if (stages < 2) return bad("resident W_hh slice leaves no room for the h ring");
    TcLstmPlan* plan = new TcLstmPlan();
    plan->smem_bytes = w_bytes + stages * H_BLK_BYTES + 1024 + 512;
    {
        cuuint64_t gdim[2] = {(cuuint64_t)H, (cuuint64_t)4 * H};
        cuuint64_t gstr[1] = {(cuuint64_t)H * 2};
        cuuint32_t box[2] = {(cuuint32_t)KBLK, (cuuint32_t)NG};
        cuuint32_t est[2] = {1, 1};
        CUresult r = g_encode_l(&plan->tmW, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(Whh_bf16), gdim, gstr, box, est,
                                CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                                CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (r != CUDA_SUCCESS) { delete plan; snprintf(msg, sizeof msg, "cuTensorMapEncodeTiled(Whh) failed: %d", (int)r); return bad(msg); }
    }
    // batch chunks: as many 128-utterance tiles per launch as fit co-resident next to the G_n unit slices
    const int tiles_per_launch = num_sms / G_n;
    const int rows_per_launch = tiles_per_launch * TILE_B;
    int chunk_index = 0;
    for (int bs = 0; bs < B; bs += rows_per_launch, ++chunk_index) {
        const int Bc = (B - bs) < rows_per_launch ? (B - bs) : rows_per_launch;
        TcLstmArgs a{};
        a.B = Bc; a.T = T; a.H = H; a.Tg = Tg; a.gx_div = gx_div; a.G_n = G_n; a.num_kb = num_kb; a.stages = stages;
        a.Gx = Gx + (size_t)bs * Tg * 4 * H;
        a.Y = reinterpret_cast<__nv_bfloat16*>(Y) + (size_t)bs * T * H;
        a.ready = ready + (size_t)(bs / TILE_B) * T;
        a.idesc = ptx::umma_idesc_f16(1, TILE_B, NG);
        CUtensorMap tm;
        cuuint64_t gdim[3] = {(cuuint64_t)H, (cuuint64_t)T, (cuuint64_t)Bc};
        cuuint64_t gstr[2] = {(cuuint64_t)H * 2, (cuuint64_t)T * H * 2};
        cuuint32_t box[3] = {(cuuint32_t)KBLK, 1, (cuuint32_t)TILE_B};
        cuuint32_t est[3] = {1, 1, 1};
        CUresult r = g_encode_l(&tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, a.Y, gdim, gstr, box, est, CU_TENSOR_MAP_INTERLEAVE_NONE,
                                CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (r != CUDA_SUCCESS) { delete plan; snprintf(msg, sizeof msg, "cuTensorMapEncodeTiled(h) failed: %d", (int)r); return bad(msg); }
        plan->tmH.push_back(tm);
        plan->args.push_back(a);
        plan->grids.push_back(dim3(G_n, ceil_div(Bc, TILE_B)));
    }
    return plan;
}

void tc_lstm_plan_destroy(TcLstmPlan* plan) { delete plan; }

int tc_lstm_ready_words(int B, int T) { return ceil_div(B, TILE_B) * T; }

cudaError_t launch_tc_lstm(const TcLstmPlan* plan, cudaStream_t s, int* launches) {
    for (size_t i = 0; i < plan->args.size(); ++i) {
        void* kargs[3] = {(void*)&plan->tmW, (void*)&plan->tmH[i], (void*)&plan->args[i]};
        // cooperative launch: the step counters make CTAs wait on one another, so co-residency must be guaranteed
        cudaError_t e = cudaLaunchCooperativeKernel((const void*)tc_lstm_kernel, plan->grids[i], dim3(192), kargs,
                                                    (size_t)plan->smem_bytes, s);
        if (e != cudaSuccess) return e;
        if (launches) ++*launches;
    }
    return cudaSuccess;
}

}  // namespace avc
