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
#include <cstdlib>
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
constexpr int NACC = 1;                   // >1: K-interleaved partial accumulators (measured: no gain, the MMA issue rate binds)
constexpr uint32_t TMEM_COLS = 64 * NACC;

struct TcLstmArgs {
    int B, T, H;
    int Tg, gx_div;              // Gx row for frame t is t / gx_div (lstm1 reads the de-duplicated projection)
    int G_n;                     // unit slices = H / 16
    int num_kb;                  // H / 64
    int stages;
    int skb;                     // K blocks (16 KiB each) per ring stage = per bulk copy
    int rotate;                  // 1: start each slice's K loop at a different block (spreads the L2 reads)
    int cs;                      // cluster size along the unit-slice axis (1 = no multicast)
    const float* Gx;             // [B][Tg][4H] fp32, unit-major gate order, bias included
    __nv_bfloat16* Y;            // [B][T][H] bf16 output sequence (also the recurrent input)
    __nv_bfloat16* Yb;           // optional recurrent copy of Y as ready-made operand tiles: [T][tiles][H/64][128][64] bf16
                                 // with the 128B swizzle already applied, so a K block is ONE contiguous 16 KiB bulk copy
    int tiles;                   // batch tiles in Yb
    int tile0;                   // first batch tile of this launch inside Yb
    unsigned int* ready;         // [batch tiles][T] step counters, zeroed before launch
    long long y_sb, y_st;        // element strides of Y along utterance / frame (batch-major: T*H, H; time-major: H, B*H)
    long long gx_sb, gx_st;      // same for Gx rows (frame index t / gx_div)
    uint32_t idesc;
    long long* trace;            // optional clock64 trace of CTA (0,0) (AVC_LSTM_TRACE=1), nullptr normally
};

constexpr int TRACE_T0 = 8, TRACE_STEPS = 8, TRACE_PTS = 8;
#define AVC_TRACE(pt) do { if (a.trace && blockIdx.x == 0 && blockIdx.y == 0 && t >= TRACE_T0 && t < TRACE_T0 + TRACE_STEPS) \
    a.trace[(t - TRACE_T0) * TRACE_PTS + (pt)] = clock64(); } while (0)

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

__device__ __forceinline__ int getenv_rot(const TcLstmArgs& a, int ns) { return (a.rotate && a.cs == 1) ? (ns * (a.num_kb / a.skb)) / a.G_n : 0; }

__global__ void __launch_bounds__(192, 1)
tc_lstm_kernel(const __grid_constant__ CUtensorMap tmW, const __grid_constant__ CUtensorMap tmH, const TcLstmArgs a) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* sW = smem;                                        // num_kb x 8 KiB, resident
    uint8_t* sH = smem + a.num_kb * W_BLK_BYTES;               // stages x 16 KiB ring
    const int stage_bytes = a.skb * H_BLK_BYTES;
    const int num_st = a.num_kb / a.skb;                       // ring stages consumed per time step
    uint64_t* bars = reinterpret_cast<uint64_t*>(sH + a.stages * stage_bytes);
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
        // a ring stage is refilled by multicast for the whole cluster, so it is free only when all cs CTAs released it
        for (int i = 0; i < a.stages; ++i) { ptx::mbar_init(&full[i], 1); ptx::mbar_init(&empty[i], (uint32_t)a.cs); }
        ptx::mbar_init(wfull, 1);
        ptx::mbar_init(acc_full, 1);
        ptx::mbar_init(acc_empty, 128);
        ptx::fence_barrier_init();
    }
    if (warp == 1) ptx::tmem_alloc<TMEM_COLS>(tmem_slot);
    ptx::tc_fence_before();
    __syncthreads();
    if (a.cs > 1) ptx::cluster_sync();     // peers' barriers are initialised before anyone multicasts / commits into them
    ptx::tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    const int kb_rot = getenv_rot(a, ns);
    const uint32_t crank = a.cs > 1 ? ptx::cluster_ctarank() : 0u;
    const uint16_t cmask = (uint16_t)((1u << a.cs) - 1u);

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
                    if (ld_acquire(ready + (t - 1)) >= (unsigned int)(4 * a.G_n)) { ok = true; break; }
                }
                if (!ok) {
                    printf("autovc_b200: lstm step counter timed out (slice %d tile %d t %d: %u/%d)\n", ns, bt, t,
                           ld_acquire(ready + (t - 1)), a.G_n);
                    __trap();
                }
                AVC_TRACE(0);
                asm volatile("fence.proxy.async.global;" ::: "memory");   // generic-proxy writes of h (other CTAs) -> TMA reads
                // rotate the K order per unit slice so that the G_n CTAs of a tile, which all start a step at the same
                // instant, do not all walk the same L2 lines of h in lock-step
                for (int ss = 0, sb = kb_rot; ss < num_st; ++ss, sb = (sb + 1 == num_st ? 0 : sb + 1)) {
                    const int kb = sb * a.skb;
                    ptx::mbar_wait(&empty[stage], phase ^ 1);
                    ptx::mbar_expect_tx(&full[stage], (uint32_t)stage_bytes);
                    if (a.Yb) {
                        // the issue of a copy costs a few hundred cycles of this single thread whatever its size
                        // (profiles/r1_lstm_step_timeline.md), so one copy brings skb K blocks
                        const __nv_bfloat16* src = a.Yb + ((((size_t)(t - 1) * a.tiles + a.tile0 + bt) * a.num_kb + kb) * TILE_B) * KBLK;
                        ptx::bulk_load(sH + stage * stage_bytes, src, (uint32_t)stage_bytes, &full[stage]);
                    } else if (a.cs == 1) {
                        ptx::tma_load_3d(sH + stage * stage_bytes, &tmH, &full[stage], kb * KBLK, b0, t - 1);
                    } else if ((uint32_t)(sb % a.cs) == crank) {
                        // every CTA of the cluster needs the same h block: fetch it from L2 once, multicast it
                        ptx::tma_load_3d_mc(sH + stage * stage_bytes, &tmH, &full[stage], kb * KBLK, b0, t - 1, cmask);
                    }
                    if (++stage == a.stages) { stage = 0; phase ^= 1; }
                }
                AVC_TRACE(1);
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
                for (int ss = 0, sb = kb_rot; ss < num_st; ++ss, sb = (sb + 1 == num_st ? 0 : sb + 1)) {
                    ptx::mbar_wait(&full[stage], phase);
                    if (ss == 0) AVC_TRACE(2);
                    ptx::tc_fence_after();
                    for (int j = 0; j < a.skb; ++j) {
                        const uint64_t adesc = ptx::umma_desc_sw128(ptx::smem_u32(sH + stage * stage_bytes + j * H_BLK_BYTES));
                        const uint64_t bdesc = ptx::umma_desc_sw128(ptx::smem_u32(sW + (sb * a.skb + j) * W_BLK_BYTES));
#pragma unroll
                        for (int k = 0; k < KBLK / 16; ++k)
                            ptx::umma_f16(tmem_base + (uint32_t)((k % NACC) * NG), adesc + (uint64_t)(k * 2), bdesc + (uint64_t)(k * 2), a.idesc,
                                          (ss != 0 || j != 0 || k >= NACC) ? 1u : 0u);
                    }
                    if (a.cs == 1) ptx::umma_commit(&empty[stage]);
                    else ptx::umma_commit_mc(&empty[stage], cmask);
                    if (++stage == a.stages) { stage = 0; phase ^= 1; }
                }
                ptx::umma_commit(acc_full);
                AVC_TRACE(3);
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
        const float* gx_row = a.Gx + (size_t)(valid ? b : 0) * a.gx_sb + (size_t)ns * NG;
        __nv_bfloat16* y_row = a.Y + (size_t)(valid ? b : 0) * a.y_sb + (size_t)ns * UNITS;

        for (int t = 0; t < a.T; ++t) {
            // prefetch the hoisted input projection for this step (independent of the recurrence)
            float4 gx[UNITS];
            const float4* gp = reinterpret_cast<const float4*>(gx_row + (size_t)(t / a.gx_div) * a.gx_st);
#pragma unroll
            for (int u = 0; u < UNITS; ++u) gx[u] = valid ? __ldg(gp + u) : make_float4(0.f, 0.f, 0.f, 0.f);

            uint32_t acc[4][16];
            if (t >= 1) {
                ptx::mbar_wait(acc_full, (uint32_t)(t - 1) & 1);
                if (et == 0) AVC_TRACE(4);
                ptx::tc_fence_after();
#pragma unroll
                for (int j = 0; j < 4; ++j) ptx::tmem_ld16(taddr + (uint32_t)(j * 16), acc[j]);
                ptx::tmem_ld_wait();
#pragma unroll
                for (int qa = 1; qa < NACC; ++qa) {      // sum the partial accumulators (each holds a K/NACC slice)
                    uint32_t part[4][16];
#pragma unroll
                    for (int j = 0; j < 4; ++j) ptx::tmem_ld16(taddr + (uint32_t)(qa * NG + j * 16), part[j]);
                    ptx::tmem_ld_wait();
#pragma unroll
                    for (int j = 0; j < 4; ++j)
#pragma unroll
                        for (int k = 0; k < 16; ++k) acc[j][k] = __float_as_uint(__uint_as_float(acc[j][k]) + __uint_as_float(part[j][k]));
                }
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
                uint4* o = reinterpret_cast<uint4*>(y_row + (size_t)t * a.y_st);
                o[0] = make_uint4(hp[0], hp[1], hp[2], hp[3]);
                o[1] = make_uint4(hp[4], hp[5], hp[6], hp[7]);
            }
            if (a.Yb) {
                // same h_t as an operand tile row: K block ns/4, 16-byte chunks (ns%4)*2 and +1, XOR-swizzled by row%8
                __nv_bfloat16* blk = a.Yb + ((((size_t)t * a.tiles + a.tile0 + bt) * a.num_kb + (ns >> 2)) * TILE_B + r) * KBLK;
                const int c0 = (ns & 3) * 2, x = r & 7;
                *reinterpret_cast<uint4*>(blk + ((c0 ^ x) << 3)) = make_uint4(hp[0], hp[1], hp[2], hp[3]);
                *reinterpret_cast<uint4*>(blk + (((c0 + 1) ^ x) << 3)) = make_uint4(hp[4], hp[5], hp[6], hp[7]);
            }
            // publish (per warp): generic-proxy stores -> visible to other CTAs' TMA (async proxy) reads, then one
            // release-increment per warp; the step counter completes at 4 * G_n
            if (et == 0) AVC_TRACE(5);
            asm volatile("fence.proxy.async.global;" ::: "memory");
            __syncwarp();
            if (et == 0) AVC_TRACE(6);
            if (lane == 0) red_release_add(ready + t, 1u);
            if (et == 0) AVC_TRACE(7);
        }
    }

    ptx::tc_fence_before();
    __syncthreads();
    if (a.cs > 1) ptx::cluster_sync();     // no CTA may exit while peers can still arrive on its barriers
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
    int cs = 1;
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

size_t tc_lstm_tile_bytes(int B, int T, int H) { return (size_t)T * ceil_div(B, TILE_B) * TILE_B * H * 2; }

TcLstmPlan* tc_lstm_plan_create(const void* Whh_bf16, const float* Gx, int Tg, int gx_div, void* Y, void* Yb, unsigned int* ready,
                                int B, int T, int H, int num_sms, bool time_major, const char** err) {
    static thread_local char msg[256];
    auto bad = [&](const char* m) -> TcLstmPlan* { if (err) *err = m; return nullptr; };
    if (!g_encode_l) return bad("tc_lstm_init() not called");
    if (H % 64 || H < 64) return bad("hidden size must be a multiple of 64");
    const int G_n = H / UNITS, num_kb = H / KBLK;
    if (G_n > num_sms) return bad("hidden size too large for one co-resident grid");
    const int w_bytes = num_kb * W_BLK_BYTES;
    const bool bulk = Yb != nullptr && !getenv("AVC_LSTM_NOBULK");
    int skb = 1;
    if (bulk) { const char* e = getenv("AVC_LSTM_SKB"); skb = e ? atoi(e) : 2; }
    while (skb > 1 && ((num_kb % skb) || (g_max_smem - 1024 - 512 - w_bytes) / (skb * H_BLK_BYTES) < 2)) skb >>= 1;
    int stages = (g_max_smem - 1024 - 512 - w_bytes) / (skb * H_BLK_BYTES);
    if (stages > MAX_STAGES) stages = MAX_STAGES;
    if (stages > num_kb / skb) stages = num_kb / skb;
    if (stages < 2) return bad("resident W_hh slice leaves no room for the h ring");
    TcLstmPlan* plan = new TcLstmPlan();
    plan->smem_bytes = w_bytes + stages * skb * H_BLK_BYTES + 1024 + 512;
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
    // cluster size for the h multicast: the largest of 8/4/2 for which one batch tile's G_n CTAs are co-resident
    int cs = 1, max_ctas = num_sms;
    const char* force = getenv("AVC_LSTM_CLUSTER");
    // measured (profiles/r1_lstm_step_timeline.md): the h stream is bound by per-SM TMA ingest, not by L2 reads, so
    // multicast buys nothing and costs co-residency; it stays available behind AVC_LSTM_CLUSTER for experiments
    for (int c = force ? atoi(force) : 1; c >= 2; c >>= 1) {
        if (G_n % c) continue;
        cudaLaunchConfig_t cfg{};
        cfg.gridDim = dim3(G_n, num_sms / G_n > 0 ? num_sms / G_n : 1);
        cfg.blockDim = dim3(192);
        cfg.dynamicSmemBytes = (size_t)plan->smem_bytes;
        cudaLaunchAttribute at[1];
        at[0].id = cudaLaunchAttributeClusterDimension;
        at[0].val.clusterDim.x = c; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
        cfg.attrs = at; cfg.numAttrs = 1;
        int ncl = 0;
        if (c > 8) cudaFuncSetAttribute(tc_lstm_kernel, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
        if (cudaOccupancyMaxActiveClusters(&ncl, tc_lstm_kernel, &cfg) == cudaSuccess && ncl * c >= G_n) {
            cs = c; max_ctas = ncl * c;
            break;
        }
        cudaGetLastError();
    }
    if (max_ctas > num_sms) max_ctas = num_sms;
    plan->cs = cs;
    const int tiles_per_launch = max_ctas / G_n;
    const int rows_per_launch = tiles_per_launch * TILE_B;
    int chunk_index = 0;
    for (int bs = 0; bs < B; bs += rows_per_launch, ++chunk_index) {
        const int Bc = (B - bs) < rows_per_launch ? (B - bs) : rows_per_launch;
        TcLstmArgs a{};
        a.B = Bc; a.T = T; a.H = H; a.Tg = Tg; a.gx_div = gx_div; a.G_n = G_n; a.num_kb = num_kb; a.stages = stages; a.skb = skb; a.cs = cs; a.rotate = getenv("AVC_LSTM_NOROT") ? 0 : 1;
        if (time_major) {   // [frame][utterance][.]: one step's rows are contiguous
            a.y_sb = H; a.y_st = (long long)B * H; a.gx_sb = 4LL * H; a.gx_st = (long long)B * 4 * H;
        } else {            // [utterance][frame][.]
            a.y_sb = (long long)T * H; a.y_st = H; a.gx_sb = (long long)Tg * 4 * H; a.gx_st = 4LL * H;
        }
        a.Gx = Gx + (size_t)bs * a.gx_sb;
        a.Y = reinterpret_cast<__nv_bfloat16*>(Y) + (size_t)bs * a.y_sb;
        a.ready = ready + (size_t)(bs / TILE_B) * T;
        a.Yb = bulk ? reinterpret_cast<__nv_bfloat16*>(Yb) : nullptr;
        a.tiles = ceil_div(B, TILE_B); a.tile0 = bs / TILE_B;
        a.idesc = ptx::umma_idesc_f16(1, TILE_B, NG);
        CUtensorMap tm;
        // dims (k, utterance, frame): the 128 rows of a box are its SECOND dimension - a box that iterates its third
        // dimension 128 times is ~4x slower in the TMA unit (measured, profiles/r1_lstm_step_timeline.md)
        cuuint64_t gdim[3] = {(cuuint64_t)H, (cuuint64_t)Bc, (cuuint64_t)T};
        cuuint64_t gstr[2] = {(cuuint64_t)a.y_sb * 2, (cuuint64_t)a.y_st * 2};
        cuuint32_t box[3] = {(cuuint32_t)KBLK, (cuuint32_t)TILE_B, 1};
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
    static long long* d_trace = nullptr;
    const bool tracing = getenv("AVC_LSTM_TRACE") != nullptr;
    if (tracing && !d_trace) cudaMalloc(&d_trace, sizeof(long long) * TRACE_STEPS * TRACE_PTS);
    for (size_t i = 0; i < plan->args.size(); ++i) {
        TcLstmArgs args = plan->args[i];
        args.trace = tracing ? d_trace : nullptr;
        if (tracing) cudaMemsetAsync(d_trace, 0, sizeof(long long) * TRACE_STEPS * TRACE_PTS, s);
        void* kargs[3] = {(void*)&plan->tmW, (void*)&plan->tmH[i], (void*)&args};
        // cooperative launch: the step counters make CTAs wait on one another, so co-residency must be guaranteed
        cudaLaunchConfig_t cfg{};
        cfg.gridDim = plan->grids[i]; cfg.blockDim = dim3(192); cfg.dynamicSmemBytes = (size_t)plan->smem_bytes; cfg.stream = s;
        cudaLaunchAttribute at[2];
        at[0].id = cudaLaunchAttributeCooperative; at[0].val.cooperative = 1;
        at[1].id = cudaLaunchAttributeClusterDimension;
        at[1].val.clusterDim.x = (unsigned)plan->cs; at[1].val.clusterDim.y = 1; at[1].val.clusterDim.z = 1;
        cfg.attrs = at; cfg.numAttrs = plan->cs > 1 ? 2 : 1;
        cudaError_t e = cudaLaunchKernelExC(&cfg, (const void*)tc_lstm_kernel, kargs);
        if (e != cudaSuccess) return e;
        if (launches) ++*launches;
        if (tracing) {   // debugging aid: synchronous dump of the per-step timeline of CTA (0,0)
            long long hbuf[TRACE_STEPS * TRACE_PTS];
            cudaStreamSynchronize(s);
            cudaMemcpy(hbuf, d_trace, sizeof hbuf, cudaMemcpyDeviceToHost);
            fprintf(stderr, "tc_lstm trace cs=%d H=%d B=%d grid=(%d,%d): per step cycles since ready-seen: tma_issued first_full mma_done acc_full_seen gates_done bar_done published | step period\n",
                    args.cs, args.H, args.B, plan->grids[i].x, plan->grids[i].y);
            for (int st = 0; st < TRACE_STEPS; ++st) {
                const long long* r = hbuf + st * TRACE_PTS;
                if (!r[0]) continue;
                fprintf(stderr, "  t=%d:", TRACE_T0 + st);
                for (int p = 1; p < TRACE_PTS; ++p) fprintf(stderr, " %6lld", r[p] - r[0]);
                if (st + 1 < TRACE_STEPS && r[TRACE_PTS]) fprintf(stderr, " | %6lld", r[TRACE_PTS] - r[0]);
                fprintf(stderr, "\n");
            }
        }
    }
    return cudaSuccess;
}

}  // namespace avc
