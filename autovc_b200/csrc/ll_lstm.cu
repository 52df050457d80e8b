// Persistent weight-stationary LSTM recurrence on the tensor cores (sm_100a).  Used for the decoder's lstm1 (H=512)
// and lstm2 (H=1024, two layers run one after the other)  [reference decoder.py:37,53,65,74].
//
// Decomposition: grid = (H/32 unit slices) x (batch tiles of 64 utterances), all CTAs co-resident (cooperative
// launch, one CTA per SM; lstm2 at B=256: 32 x 4 = 128 CTAs).  The GEMM of a step is computed TRANSPOSED:
//   D[128 gate rows x 64 utterances] (TMEM, fp32) = W_slice[128 x H] . h_{t-1}^T[H x 64]
// so that the resident operand is A and the streamed operand is B:
//   * A = the CTA's 128 gate rows (32 hidden units x i,f,g,o, unit-major packing) of W_hh.  They live in TENSOR
//     MEMORY for the whole sequence (448 of the 512 columns = K 0..895; the rest of K, if any, in shared memory),
//     so an MMA reads only the 2 KiB h block from shared memory and runs at the tensor-pipe floor.
//   * B = h_{t-1} of the CTA's 64 utterances, [64 x H] bf16 = 128 KiB per step at H=1024 - half of what a
//     128-utterance x 16-unit tiling must ingest per SM and per step.  The measured bound of this kernel family is
//     the L2 -> SM path (latency 1.2-2k cycles, ~4 KB/clk over the whole chip; profiles/r1_lstm_step_timeline.md),
//     so bytes per SM and round trips per step are what the design minimises.
//
// Exchange protocol ("the data is the flag"): CTAs exchange h through an L2-resident buffer Yx laid out as
// [T][batch tile][H/8 planes][64 rows][8 bf16], pre-filled with the bit pattern 0xFFFF (a NaN that the float->bf16
// conversion never produces: it canonicalises NaN to 0x7FFF).  Producers simply store their 16-byte pieces of h_t;
// consumers read with strong (L1-bypassing) coalesced 16-byte loads and treat any 32-bit word of all ones as "not
// written yet" (32-bit accesses are single-copy atomic) and poll it again.  No step counter, no release/acquire fence
// pair, no second round trip between "flag seen" and "data fetched".  All of a step's loads of a warp are in flight
// together (one wave), and a warp owns one K chunk (two producer CTAs), so chunks are consumed as their producers
// finish and the skew between CTAs overlaps with the stream.
//
// Warp roles (512 threads; 128 registers per thread):
//   warps 0-15  loaders : one K chunk (or part of one) each: poll a 16-byte canary per lane, then fetch the chunk's
//                         pieces, validate, store them as a 128B-swizzled K-major operand tile (generic-proxy stores +
//                         fence.proxy.async) and arrive on the chunk's mbarrier
//   warps 0-7   epilogue: tcgen05.ld the gate pre-activations (lane = gate row, 32 utterance columns per thread), add the
//                         hoisted input projection Gx_t (staged in shared memory with cp.async), activation, 4x4 quad
//                         transpose with shuffles (lane j ends with i,f,g,o of 8 utterances), cell update (c in registers),
//                         h_t -> shared-memory transpose -> 16-byte stores to Yx (exchange) and Y (next layer)
//   MMA issue           : by warp 15 after its own chunk, whole-warp with uniform operands (one elected lane), chunks in a
//                         fixed order (bitwise reproducible accumulation), gated by one mbarrier per group of four loader warps
// The input projection x_t W_ih^T + b_ih + b_hh for all t is hoisted into one dense GEMM (tc_gemm.cu) beforehand.
//
// Every spin/wait is bounded: a protocol bug traps (CUDA error) instead of hanging the GPU.
#include <cstdio>
#include <cstdlib>
#include <type_traits>
#include <utility>
#include <vector>
#include <cuda.h>

#include "common.cuh"
#include "launch.h"
#include "ptx.cuh"

namespace avc {

namespace {

constexpr int SLICE_UNITS = 32;           // hidden units per CTA
constexpr int MROWS = 4 * SLICE_UNITS;    // gate rows per CTA = UMMA M = TMEM lanes
constexpr int TILE_N = 64;                // utterances per CTA PAIR = UMMA N (each CTA stages half of the rows)
constexpr int ROWS_CTA = TILE_N / 2;      // utterance rows of h staged by one CTA
constexpr int KBLK = 64;                  // K elements per chunk (128 bytes per operand row)
constexpr int W_BLK_BYTES = MROWS * KBLK * 2;     // 16 KiB (shared-memory part of W)
constexpr int H_BLK_BYTES = ROWS_CTA * KBLK * 2;  // 4 KiB per CTA and chunk
constexpr int PLANE_ELEMS = 8;            // Yx plane width: 8 bf16 = 16 B
constexpr int PLANE_STRIDE = TILE_N * PLANE_ELEMS;      // elements between planes
constexpr int LOAD_WARPS = 16, EPI_WARPS = 8;
constexpr int MMA_WARP0 = LOAD_WARPS - 2;  // the last two loader warps of the LEADER CTA also issue the MMAs (one K half each)
constexpr int NUM_THREADS = LOAD_WARPS * 32;            // 512: four warps per scheduler, 128 registers per thread
constexpr int EPI_THREADS = EPI_WARPS * 32;             // 256
constexpr int GX_BYTES = TILE_N * MROWS * 4;            // 32 KiB: Gx_t of the tile, [utterance][gate row] fp32
constexpr int HT_BYTES = TILE_N * SLICE_UNITS * 2;      // 4 KiB: h_t transpose staging, [utterance][unit] bf16
constexpr uint32_t D_COLS = 2 * TILE_N;   // two accumulators (one per issuing warp: K halves), summed in a fixed order
constexpr uint32_t W_COL0 = D_COLS;       // W chunks follow: 32 columns each
constexpr int MAX_TMEM_CHUNKS = (512 - (int)D_COLS) / 32;   // 12
constexpr uint32_t TMEM_COLS = 512;
constexpr uint32_t SENTINEL = 0xFFFFFFFFu;
constexpr int MAX_KB = 16;

struct LlLstmArgs {
    int B, T, H;
    int gx_div;                  // Gx row for frame t is t / gx_div (lstm1 reads the de-duplicated projection)
    const void* Gx;              // [B][Tg][4H] fp32 (or fp16: GXH kernels), unit-major gate order, bias included
    const __nv_bfloat16* Whh;    // [4H][H] bf16, unit-major gate rows
    __nv_bfloat16* Y;            // [B][T][H] bf16 output sequence for the next layer
    __nv_bfloat16* Yx;           // exchange copy: [T][tiles][H/8][64][8] bf16, pre-filled with 0xFF
    int tiles, tile0;            // batch tiles in Yx, first tile of this launch
    long long y_sb, y_st;        // element strides of Y along utterance / frame
    long long gx_sb, gx_st;      // same for Gx rows (frame index t / gx_div)
    uint32_t idesc;
    int approx;                  // 1 (default): gate activations with MUFU.TANH (tanh.approx.f32, abs err ~5e-4: below the bf16 rounding of h;
                                 // measured end-to-end error unchanged); AVC_LSTM_EXACT=1 selects ex2/rcp based sigmoid/tanh
    int canary;                  // AVC_LSTM_CANARY: 0 no canary poll, 1 (default) loader-only warps, 2 all warps (the previous protocol)
    int gx_late;                 // 1 (default): issue the Gx_t staging after the operand tile is published - fence.proxy.async waits for the
                                 // warp's outstanding copies, and Gx_t comes from HBM (lstm2: 1.29 -> 1.24 ms); AVC_LSTM_GX_LATE=0: behind the h loads
    long long* trace;            // optional clock64 trace of CTA (0,0) (AVC_LSTM_TRACE=1), nullptr normally
};

constexpr int TRACE_T0 = 8, TRACE_STEPS = 8, TRACE_PTS = 32;
constexpr int TRACE_WORDS = (2 * TRACE_STEPS + 1) * TRACE_PTS;   // 0..5 phase stamps, 8+kc: chunk kc seen by the MMA thread
#define AVC_TRACE(pt) do { if (a.trace && blockIdx.x == 0 && blockIdx.y == 0 && t >= TRACE_T0 && t < TRACE_T0 + TRACE_STEPS) \
    a.trace[(t - TRACE_T0) * TRACE_PTS + (pt)] = clock64(); } while (0)

__device__ __forceinline__ uint4 ld_strong(const void* p) {
    uint4 v;
    asm volatile("ld.relaxed.gpu.global.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_strong(void* p, uint4 v) {
    asm volatile("st.relaxed.gpu.global.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
__device__ __forceinline__ bool unwritten(const uint4& v) {
    return v.x == SENTINEL || v.y == SENTINEL || v.z == SENTINEL || v.w == SENTINEL;
}
__device__ __forceinline__ void st_shared_v4(uint32_t addr, uint4 v) {
    asm volatile("st.shared.v4.u32 [%0], {%1,%2,%3,%4};" ::"r"(addr), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
__device__ __forceinline__ uint4 ld_shared_v4(uint32_t addr) {
    uint4 v;
    asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(addr) : "memory");
    return v;
}
__device__ __forceinline__ float ld_shared_f32(uint32_t addr) {
    float v;
    asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(addr) : "memory");
    return v;
}
__device__ __forceinline__ void st_shared_u16(uint32_t addr, uint16_t v) {
    asm volatile("st.shared.u16 [%0], %1;" ::"r"(addr), "h"(v) : "memory");
}
__device__ __forceinline__ void cp_async16(uint32_t dst_smem, const void* src) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst_smem), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_group 0;" ::: "memory"); }
// c' = f c + i g with ONE fixed rounding order in every kernel of this file (rows of a batch must not depend on which kernel
// variant the batch size selects): the product i g is rounded, then fused into f c
__device__ __forceinline__ float cell_update(float fg, float c, float ig, float gg) { return fmaf(fg, c, __fmul_rn(ig, gg)); }
__device__ __forceinline__ void named_bar_sync(int id, int threads) { asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(threads) : "memory"); }

template <int NKB, bool GXH, int TN>
__global__ void __launch_bounds__(NUM_THREADS, 1)
ll_lstm_kernel(const __grid_constant__ CUtensorMap tmW, const LlLstmArgs a) {
    constexpr int NKB_T = NKB < MAX_TMEM_CHUNKS ? NKB : MAX_TMEM_CHUNKS;   // W chunks resident in tensor memory
    constexpr int NKB_S = NKB - NKB_T;                                     // W chunks resident in shared memory
    constexpr int WPC = LOAD_WARPS / NKB;                                  // loader warps per chunk: 1, 2 or 4
    // TN = utterances per CTA pair = UMMA N.  64 fills the machine at B = 256; 32 is used when the batch leaves room for twice
    // the tiles: a step is a latency chain, and half the tile halves its MMA, load and epilogue stretches.
    constexpr int ROWS = TN / 2;                          // utterance rows of h staged by one CTA of the pair
    constexpr int HBLK = ROWS * KBLK * 2;                 // bytes per CTA and chunk of the operand tile
    constexpr int PSTRIDE = TN * PLANE_ELEMS;             // elements between planes of Yx
    constexpr int CPT = TN / 2;                           // accumulator columns (utterances) per epilogue thread
    static_assert(TN == 64 || TN == 32, "tile of 64 or 32 utterances");
    static_assert(NKB == 16 || NKB == 8 || NKB == 4, "H must be 1024, 512 or 256");

    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* sW = smem;                                        // NKB_S x 16 KiB: W columns beyond the TMEM-resident part
    uint8_t* sH = sW + NKB_S * W_BLK_BYTES;                    // NKB x 4 KiB: this CTA's 32 rows of the h_{t-1} operand tiles
    uint8_t* sGx = sH + NKB * HBLK;                     // 32 KiB
    uint8_t* sHT = sGx + GX_BYTES;                             // 4 KiB
    uint64_t* bars = reinterpret_cast<uint64_t*>(sHT + HT_BYTES);
    uint64_t* wfull = bars;
    uint64_t* gfull = bars + 4;                                // [4] the chunks of loader-warp group g (warps 4g..4g+3) are in shared memory
    uint64_t* acc_full = wfull + 1;                            // the MMAs of a step are complete
    uint64_t* acc_empty = wfull + 2;                           // the epilogue has read the accumulator
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(wfull + 3);

    const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0), lane = threadIdx.x & 31;   // warp: provably uniform
    const int ns = blockIdx.x;            // unit slice (32 units)
    const int bt = blockIdx.y;            // batch tile (64 utterances)
    const int b0 = bt * TN;
    const uint32_t crank = ptx::cluster_ctarank();   // 0 = leader of the CTA pair (adjacent unit slices, same batch tile)
    const size_t tile_elems = (size_t)a.H * TN;                         // one (t, tile) block of Yx
    const size_t step_elems = tile_elems * a.tiles;
    __nv_bfloat16* yx_tile = a.Yx + (size_t)(a.tile0 + bt) * tile_elems;    // + t * step_elems

    if (warp == 0) {
        if (lane == 0) {
            if (NKB_S > 0) ptx::prefetch_tmap(&tmW);
            ptx::mbar_init(wfull, 1);
            for (int g = 0; g < 4; ++g) ptx::mbar_init(&gfull[g], 8);     // 4 loader warps of each CTA of the pair
            ptx::mbar_init(acc_full, 2);                                  // one multicast commit per issuing warp
            ptx::mbar_init(acc_empty, 2 * EPI_WARPS);                     // epilogue warps of both CTAs
            ptx::fence_barrier_init();
        }
        __syncwarp();
        ptx::tmem_alloc_pair<TMEM_COLS>(tmem_slot);
    }
    ptx::tc_fence_before();
    __syncthreads();
    ptx::cluster_sync();     // the peer's barriers and tensor memory are set up before anyone arrives on them / issues pair MMAs
    ptx::tc_fence_after();
    // the CTA owns the SM and allocates all 512 columns, so the allocation starts at column 0 / lane 0: treating the base as
    // the constant 0 lets the compiler keep every TMEM address and descriptor of the MMA loop in uniform registers
    if (*tmem_slot != 0u) { if (threadIdx.x == 0) printf("autovc_b200: unexpected TMEM base %u\n", *tmem_slot); __trap(); }
    constexpr uint32_t tmem_base = 0u;

    // ---- resident W: chunks 0..NKB_T-1 -> tensor memory (lane = gate row, 32 columns per chunk), the rest -> smem ----
    if (warp < 4) {
        const int R = warp * 32 + lane;
        const __nv_bfloat16* wrow = a.Whh + (size_t)(ns * MROWS + R) * a.H;
        const uint32_t wl = tmem_base + W_COL0 + ((uint32_t)(warp * 32) << 16);
        for (int kc = 0; kc < NKB_T; ++kc) {
            uint4 v[8];
#pragma unroll
            for (int i = 0; i < 8; ++i) v[i] = __ldg(reinterpret_cast<const uint4*>(wrow + kc * KBLK) + i);
            ptx::tmem_st32(wl + (uint32_t)(kc * 32), reinterpret_cast<const uint32_t(&)[32]>(v));
        }
        ptx::tmem_st_wait();
    }
    if (NKB_S > 0 && warp == 4 && lane == 0) {
        ptx::mbar_expect_tx(wfull, (uint32_t)(NKB_S * W_BLK_BYTES));
        for (int i = 0; i < NKB_S; ++i)
            ptx::tma_load_2d(sW + i * W_BLK_BYTES, &tmW, wfull, (NKB_T + i) * KBLK, ns * MROWS);
        ptx::mbar_wait(wfull, 0);
    }
    ptx::tc_fence_before();
    __syncthreads();
    ptx::cluster_sync();     // both CTAs' W is resident before the leader issues
    ptx::tc_fence_after();

    {
        // ===================== loaders (warps 0..15) + epilogue (warps 0..7) =====================
        const bool epi = warp < EPI_WARPS;
        // ---- loader geometry: this warp's chunk and, per piece i, the (row, plane) of the lane ----
        const int kc = warp / WPC, sub = warp % WPC;
        constexpr int NP = ROWS * 8 * NKB / NUM_THREADS;      // 16-byte pieces per lane and step: ROWS rows x 8 planes x NKB chunks / 512
        auto piece_row = [&](int i) {
            if constexpr (TN == 64) return WPC == 1 ? lane : (WPC == 2 ? sub * 16 + (lane & 15) : sub * 8 + (lane & 7));
            else return lane & 15;
        };
        auto piece_plane = [&](int i) {
            if constexpr (TN == 64) return WPC == 1 ? i : (WPC == 2 ? 2 * i + (lane >> 4) : 4 * i + (lane >> 3));
            else return sub * (8 / WPC) + 2 * i + (lane >> 4);
        };
        const uint32_t sH_chunk = ptx::smem_u32(sH) + (uint32_t)(kc * HBLK);
        const __nv_bfloat16* chunk_base = yx_tile + (size_t)kc * 8 * PSTRIDE + (size_t)(crank * ROWS) * PLANE_ELEMS;

        // ---- epilogue geometry ----
        const int q = warp & 3, half = (warp >> 2) & 1;       // TMEM lane quarter; utterance half (CPT columns)
        const int R = q * 32 + lane;                          // gate row of this CTA = TMEM lane
        const int gj = lane & 3;                              // gate of this lane: 0=i 1=f 2=g 3=o
        const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(half * CPT);
        const float act_s = gj == 2 ? 2.0f : 1.0f, act_o = gj == 2 ? -1.0f : 0.0f;   // tanh(x) = 2 sigmoid(2x) - 1
        // the sigmoid gates' rows of W_ih, W_hh and the bias are packed pre-multiplied by 0.5 (api.cu::pack_lstm): the
        // pre-activation x' is 0.5 x for i, f, o and x for g, i.e. always the argument of the tanh below
        float c[CPT / 4];
#pragma unroll
        for (int k = 0; k < CPT / 4; ++k) c[k] = 0.0f;
        const uint32_t gx_smem = ptx::smem_u32(sGx);
        const uint32_t ht_smem = ptx::smem_u32(sHT);
        // Gx staging: thread e copies 16-byte pieces e + 256 i of the tile's [utterance][gate row] block
        auto stage_gx = [&](int t) {
            const size_t toff = (size_t)(t / a.gx_div) * a.gx_st + (size_t)ns * MROWS;
            constexpr int ES = GXH ? 2 : 4, SEGS = MROWS * ES / 16;       // 16-byte pieces per utterance: 16 (fp16) or 32 (fp32)
#pragma unroll
            for (int i = 0; i < TN * SEGS / EPI_THREADS; ++i) {
                const int id = threadIdx.x + EPI_THREADS * i;
                const int u = id / SEGS, seg = id % SEGS;
                const int b = b0 + u < a.B ? b0 + u : a.B - 1;
                cp_async16(gx_smem + (uint32_t)(u * (MROWS * ES) + seg * 16),
                           reinterpret_cast<const char*>(a.Gx) + ((size_t)b * a.gx_sb + toff) * ES + seg * 16);
            }
            cp_async_commit();
        };
        auto gx_at = [&](int k) -> float {      // Gx_t[utterance half*32 + k][gate row R]
            if constexpr (GXH) {
                uint16_t hbits;
                asm volatile("ld.shared.u16 %0, [%1];" : "=h"(hbits) : "r"(gx_smem + (uint32_t)(((half * CPT + k) * MROWS + R) * 2)));
                return __half2float(__ushort_as_half(hbits));
            } else {
                return ld_shared_f32(gx_smem + (uint32_t)(((half * CPT + k) * MROWS + R) * 4));
            }
        };
        // output geometry of the final 16-byte stores: threads e < 4 TN -> utterance e/4, 8-unit piece e%4
        const int ou = (threadIdx.x >> 2) & (TN - 1), op = threadIdx.x & 3;
        const bool ostore = threadIdx.x < TN * 4;
        const bool ovalid = ostore && b0 + ou < a.B;
        __nv_bfloat16* y_out = a.Y + (size_t)(ovalid ? b0 + ou : 0) * a.y_sb + (size_t)ns * SLICE_UNITS + op * 8;
        __nv_bfloat16* yx_out = yx_tile + ((size_t)(ns * 4 + op) * TN + ou) * PLANE_ELEMS;
        if (epi) stage_gx(0);

        for (int t = 0; t < a.T; ++t) {
            if (t >= 1) {
                // ---- h_{t-1} chunk: L2 -> registers (poll) -> swizzled operand tile in shared memory ----
                const __nv_bfloat16* src = chunk_base + (size_t)(t - 1) * step_elems;
                // Epilogue warps get here right after storing their own h_{t-1}: the other slices' pieces are then (measured) already
                // visible, so they fetch the whole step in one wave and re-fetch only what is still missing - a separate poll before
                // the wave would be a second L2 round trip (~2k cycles under load) per step.  Loader-only warps arrive a whole epilogue
                // early; a single 16-byte canary per lane keeps them off the fabric until their producers have stored.
                if (epi ? a.canary >= 2 : a.canary >= 1) {
                    const __nv_bfloat16* p = src + (size_t)piece_plane(0) * PSTRIDE + piece_row(0) * PLANE_ELEMS;
                    uint4 cn = ld_strong(p);
                    for (uint32_t spin = 0; unwritten(cn); ++spin) {
                        cn = ld_strong(p);
                        if (spin > (1u << 22)) { printf("autovc_b200: lstm h canary timed out (slice %d tile %d t %d warp %d)\n", ns, bt, t, warp); __trap(); }
                    }
                }
                if (threadIdx.x == 0) AVC_TRACE(0);
                uint4 v[NP];
#pragma unroll
                for (int i = 0; i < NP; ++i)
                    v[i] = ld_strong(src + (size_t)piece_plane(i) * PSTRIDE + piece_row(i) * PLANE_ELEMS);
                if (epi && !a.gx_late) stage_gx(t);    // (A/B knob) Gx_t behind the h loads
                // the MMAs of step t-1 still read the tiles until acc_full(t-1): normally long past
                if (t >= 2 && !epi) ptx::mbar_wait(acc_full, (uint32_t)(t - 2) & 1);
                for (uint32_t spin = 0;; ++spin) {          // pieces not written yet: fetch all of them again, together
                    bool missing = false;
#pragma unroll
                    for (int i = 0; i < NP; ++i)
                        if (unwritten(v[i])) {
                            v[i] = ld_strong(src + (size_t)piece_plane(i) * PSTRIDE + piece_row(i) * PLANE_ELEMS);
                            missing = true;
                        }
                    if (!missing) break;
                    if (spin > (1u << 22)) {
                        printf("autovc_b200: lstm h poll timed out (slice %d tile %d t %d chunk %d lane %d)\n", ns, bt, t, kc, lane);
                        __trap();
                    }
                }
                if (threadIdx.x == 0) AVC_TRACE(9);
#pragma unroll
                for (int i = 0; i < NP; ++i) {
                    const int row = piece_row(i), pl = piece_plane(i);
                    st_shared_v4(sH_chunk + (uint32_t)(row * 128 + ((pl ^ (row & 7)) << 4)), v[i]);
                }
                ptx::fence_proxy_async_smem();          // generic-proxy smem writes -> tcgen05.mma (async proxy) reads
                __syncwarp();
                if (lane == 0) ptx::mbar_arrive_cluster(&gfull[warp >> 2], 0u);     // the LEADER's barrier (both CTAs arrive)
                if (epi && a.gx_late) stage_gx(t);     // needed a whole MMA phase from now; issued after the fence above, which would wait for it
                if (crank == 0 && warp >= MMA_WARP0) {
                    // Leader CTA, warps 14/15: issue the pair's MMAs of the step, one K half each into its own accumulator,
                    // chunk by chunk in a fixed order (the accumulation order must not depend on timing: results are bitwise
                    // reproducible).  Whole-warp, uniform operands, one elected lane issues.
                    const int ih = warp - MMA_WARP0;
                    if (t >= 2) ptx::mbar_wait(acc_empty, (uint32_t)(t - 2) & 1);   // both epilogues of step t-1 have read D
                    // instantiated per issuing warp: every chunk index, TMEM address and descriptor offset is then a compile-time
                    // constant (a runtime K half doubles the code with TS/SS branches and puts ~20 instructions between two MMAs)
                    auto issue = [&](auto IH) {
                        constexpr int ih_ = decltype(IH)::value;
                        constexpr uint32_t d_tmem = tmem_base + (uint32_t)(ih_ * TILE_N);
                        const uint32_t sH_u = ptx::smem_u32(sH), sW_u = ptx::smem_u32(sW);
                        auto issue_group = [&](auto G) {
                            constexpr int g = decltype(G)::value;
                            ptx::mbar_wait(&gfull[g], (uint32_t)(t - 1) & 1);
                            ptx::tc_fence_after();
                            if (g == 0) AVC_TRACE(2);
                            AVC_TRACE(10 + g);
                            [&]<int... CC>(std::integer_sequence<int, CC...>) {
                                ([&] {
                                    constexpr int c = g * (NKB / 4) + CC;
                                    constexpr uint32_t first = (g == ih_ * 2 && CC == 0) ? 0u : 1u;
                                    const uint64_t bdesc = ptx::umma_desc_sw128(sH_u + (uint32_t)(c * HBLK));
                                    // N=32 MMAs retire every 17 cycles: one block per chunk (one election, immediates) keeps the issue
                                    // ahead of the pipe; at N=64 (33 cycles) separate MMAs measured faster (operand set-up overlaps)
                                    if constexpr (c < NKB_T) {
                                        constexpr uint32_t a_tmem = tmem_base + W_COL0 + (uint32_t)(c * 32);
                                        if constexpr (TN == 32) {
                                            ptx::umma2_f16_ts_chunk_warp<d_tmem, a_tmem>((uint32_t)bdesc, (uint32_t)(bdesc >> 32), a.idesc, first);
                                        } else {
#pragma unroll
                                            for (int k = 0; k < KBLK / 16; ++k)
                                                ptx::umma2_f16_ts_warp(d_tmem, a_tmem + (uint32_t)(k * 8), bdesc + (uint64_t)(k * 2), a.idesc, k != 0 ? 1u : first);
                                        }
                                    } else {
                                        const uint64_t adesc = ptx::umma_desc_sw128(sW_u + (uint32_t)((c - NKB_T) * W_BLK_BYTES));
                                        if constexpr (TN == 32) {
                                            ptx::umma2_f16_ss_chunk_warp<d_tmem>((uint32_t)adesc, (uint32_t)(adesc >> 32), (uint32_t)bdesc,
                                                                                 (uint32_t)(bdesc >> 32), a.idesc, first);
                                        } else {
#pragma unroll
                                            for (int k = 0; k < KBLK / 16; ++k)
                                                ptx::umma2_f16_warp(d_tmem, adesc + (uint64_t)(k * 2), bdesc + (uint64_t)(k * 2), a.idesc, k != 0 ? 1u : first);
                                        }
                                    }
                                }(), ...);
                            }(std::make_integer_sequence<int, NKB / 4>{});
                        };
                        issue_group(std::integral_constant<int, ih_ * 2>{});
                        issue_group(std::integral_constant<int, ih_ * 2 + 1>{});
                    };
                    if (ih == 0) issue(std::integral_constant<int, 0>{}); else issue(std::integral_constant<int, 1>{});
                    ptx::umma2_commit_warp(acc_full);       // -> acc_full of BOTH CTAs
                    if (ih == 1) AVC_TRACE(3); else AVC_TRACE(14);
                }
                if (threadIdx.x == 0) AVC_TRACE(1);
            }
            if (!epi) continue;

            // ---- epilogue: gates, cell state, h_t ----
            cp_async_wait_all();                       // own pieces of Gx_t ...
            named_bar_sync(2, EPI_THREADS);            // ... and everybody else's
            float act[CPT];
            if (t >= 1) {
                ptx::mbar_wait(acc_full, (uint32_t)(t - 1) & 1);
                if (threadIdx.x == 0) AVC_TRACE(4);
                ptx::tc_fence_after();
                uint32_t acc0[CPT / 16][16], acc1[CPT / 16][16];
#pragma unroll
                for (int j = 0; j < CPT / 16; ++j) {
                    ptx::tmem_ld16(taddr + (uint32_t)(16 * j), acc0[j]);
                    ptx::tmem_ld16(taddr + (uint32_t)(TILE_N + 16 * j), acc1[j]);
                }
                ptx::tmem_ld_wait();
                ptx::tc_fence_before();
                __syncwarp();
                if (lane == 0) ptx::mbar_arrive_cluster(acc_empty, 0u);
#pragma unroll
                for (int k = 0; k < CPT; ++k) act[k] = __uint_as_float(acc0[k >> 4][k & 15]) + __uint_as_float(acc1[k >> 4][k & 15]);
            } else {
#pragma unroll
                for (int k = 0; k < CPT; ++k) act[k] = 0.0f;
            }
            // + Gx_t[utterance half*CPT + k][gate row R], then this lane's activation
            if (a.approx) {
                // sigmoid(x) = 0.5 tanh(0.5 x) + 0.5, tanh(x) as is: one MUFU.TANH per gate value
                const float ps = gj == 2 ? 1.0f : 0.5f, po = gj == 2 ? 0.0f : 0.5f;
#pragma unroll
                for (int k = 0; k < CPT; ++k) {
                    const float x = act[k] + gx_at(k);
                    float th;
                    asm("tanh.approx.f32 %0, %1;" : "=f"(th) : "f"(x));
                    act[k] = fmaf(ps, th, po);
                }
            } else {
#pragma unroll
                for (int k = 0; k < CPT; ++k) {
                    const float x = act[k] + gx_at(k);
                    act[k] = __fdividef(act_s, 1.0f + __expf(-2.0f * x)) + act_o;
                }
            }
            if (threadIdx.x == 0) AVC_TRACE(6);
            // 4x4 quad transpose (blocks of CPT/4 columns): lane j ends with all four gates of columns (CPT/4) j .. (CPT/4) j + CPT/4 - 1
            constexpr int H1 = CPT / 2, H2 = CPT / 4;
            float X[H1], Y2[H1];
#pragma unroll
            for (int k = 0; k < H1; ++k) {
                const float send = (gj & 2) ? act[k] : act[H1 + k];
                X[k] = (gj & 2) ? act[H1 + k] : act[k];                      // own gate gj
                Y2[k] = __shfl_xor_sync(0xffffffffu, send, 2);               // gate gj ^ 2
            }
            uint32_t hp[H2 / 2];
#pragma unroll
            for (int k = 0; k < H2; k += 2) {
                float hv[2];
#pragma unroll
                for (int w = 0; w < 2; ++w) {
                    const int kk = k + w;
                    const float sx = (gj & 1) ? X[kk] : X[H2 + kk], sy = (gj & 1) ? Y2[kk] : Y2[H2 + kk];
                    const float kx = (gj & 1) ? X[H2 + kk] : X[kk], ky = (gj & 1) ? Y2[H2 + kk] : Y2[kk];
                    const float rx = __shfl_xor_sync(0xffffffffu, sx, 1), ry = __shfl_xor_sync(0xffffffffu, sy, 1);
                    // kx: gate gj, ky: gj^2, rx: gj^1, ry: gj^3
                    const float e0 = (gj & 1) ? rx : kx, e2 = (gj & 1) ? ry : ky;     // gates (gj & 2) and (gj & 2) ^ 2
                    const float e1 = (gj & 1) ? kx : rx, e3 = (gj & 1) ? ky : ry;     // gates (gj & 2) + 1 and its ^ 2
                    const float ig = (gj & 2) ? e2 : e0, gg = (gj & 2) ? e0 : e2;
                    const float fg = (gj & 2) ? e3 : e1, og = (gj & 2) ? e1 : e3;
                    c[kk] = cell_update(fg, c[kk], ig, gg);
                    float tc;
                    if (a.approx) asm("tanh.approx.f32 %0, %1;" : "=f"(tc) : "f"(c[kk]));
                    else tc = 1.0f - __fdividef(2.0f, 1.0f + __expf(2.0f * c[kk]));
                    hv[w] = og * tc;
                }
                __nv_bfloat162 h2 = __floats2bfloat162_rn(hv[0], hv[1]);
                hp[k >> 1] = *reinterpret_cast<uint32_t*>(&h2);
            }
            if (threadIdx.x == 0) AVC_TRACE(7);
            // transpose through shared memory: hT[utterance][unit], 16-byte pieces XOR-swizzled by (utterance / 8) & 3
            {
                const int unit = q * 8 + (lane >> 2);
#pragma unroll
                for (int k = 0; k < H2; ++k) {
                    const int u = half * CPT + gj * H2 + k;
                    const int piece = (unit >> 3) ^ ((u >> 3) & 3);
                    const uint16_t hv16 = (uint16_t)((k & 1) ? (hp[k >> 1] >> 16) : (hp[k >> 1] & 0xffffu));
                    st_shared_u16(ht_smem + (uint32_t)(u * 64 + piece * 16 + (unit & 7) * 2), hv16);
                }
            }
            named_bar_sync(1, EPI_THREADS);
            if (threadIdx.x == 0) AVC_TRACE(8);
            if (ostore) {
                const uint4 hv4 = ld_shared_v4(ht_smem + (uint32_t)(ou * 64 + ((op ^ ((ou >> 3) & 3)) << 4)));
                // utterances beyond B hold bounded garbage: they are exchanged like any other row, so loaders need no
                // row masks, and they are never written to Y
                if (t + 1 < a.T) st_strong(yx_out + (size_t)t * step_elems, hv4);
                if (ovalid) *reinterpret_cast<uint4*>(y_out + (size_t)t * a.y_st) = hv4;
            }
            if (threadIdx.x == 0) AVC_TRACE(5);
        }
    }

    ptx::tc_fence_before();
    __syncthreads();
    ptx::cluster_sync();     // no CTA may exit (or free tensor memory) while its peer can still arrive on its barriers
    if (warp == 0) {
        ptx::tc_fence_after();
        ptx::tmem_dealloc_pair<TMEM_COLS>(tmem_base);
    }
}

// =====================================================================================================================
// Pipelined variant: TWO independent 32-utterance sub-tiles per CTA pair.
//
// A step of one tile is a dependent chain  exchange (store -> visible -> load wave) -> MMAs -> gate epilogue  in which every
// unit idles while the others work.  Two sub-tiles of one CTA pair are independent chains over the same resident W_hh: while
// sub-tile A's h crosses L2, sub-tile B's MMAs and epilogue run.  Roles (17 warps, 96 registers per thread):
//   warps 0-7    sub-tile 0: loaders of its h_{t-1} rows (16 rows per CTA x H) AND its gate epilogue (16 columns per thread)
//   warps 8-15   sub-tile 1: the same
//   warps 16,17  MMA issuers (leader CTA), one K half each into its own accumulator (the one-tile kernel's split, so a row's
//                result is bitwise the same whichever kernel its batch size selects), each strictly alternating
//                A(t), B(t), A(t+1), ... in program order.  That order (a) keeps the accumulation order independent of
//                timing and (b) pins the two chains one MMA phase apart: two free-running chains fall into lockstep and then
//                share every unit at half rate, which is the one-tile kernel again (profiles/r1_ll_lstm_step_timeline.md).
//                One warp issues an N = 32 MMA every ~27 cycles (descriptor set-up), the pipe retires one every ~17: two.
// Two accumulators per sub-tile (128 columns in all) leave room for 12 of the 16 K chunks of W_hh in tensor memory.
// The 16 group warps never meet at a CTA-level barrier inside the time loop: a warp waits for its own h pieces, for the
// accumulator mbarrier and for nothing else.  (Profile of the barrier version, profiles/r2_ll2_lstm_notes.md: 15 % of all warp
// samples sat at the group barrier in front of the epilogue, i.e. behind the group's slowest poller, and another 5 % at the one in
// front of the stores.)  Hence: the Gx_t block is staged warp-privately (a warp copies what its own lanes read), the 4x4 gate
// transpose goes through a per-warp shared-memory tile (16 STS + 4 LDS.128 instead of 16 shuffles + ~60 selects), h_t leaves
// straight from registers as 32-bit words (one lane-pair exchange), and the sigmoid rows arrive pre-scaled
// (api.cu::pack_lstm).  Exchange protocol and operand layout are the one-tile kernel's.  No acc_empty barrier: a
// sub-tile's operands for step t+1 are published by the very warps that read its accumulators of step t (after their named
// barrier), so "operands complete" implies "accumulators drained" in both CTAs.
// =====================================================================================================================
namespace p2 {
constexpr int SUB_N = 32;                            // utterances per sub-tile = UMMA N
constexpr int SUB_ROWS = SUB_N / 2;                  // rows of h staged per CTA and sub-tile
constexpr int GROUP_WARPS = 8, GROUP_THREADS = GROUP_WARPS * 32;
constexpr int ISSUE_WARP = 2 * GROUP_WARPS;          // 16, 17: one K half each
constexpr int THREADS = (2 * GROUP_WARPS + 2) * 32;  // 576
constexpr uint32_t D_COLS2 = 4 * SUB_N;              // two accumulators (K halves) per sub-tile: columns (2 g + ih) * 32
constexpr int MAX_TMEM_CHUNKS2 = (512 - (int)D_COLS2) / 32;    // 12
constexpr int GT_BYTES = GROUP_WARPS * 16 * 32 * 4;  // 16 KiB per sub-tile: per warp a [16 columns][32 gate rows] fp32 transpose tile
constexpr int HBLK2 = SUB_ROWS * KBLK * 2;           // 2 KiB per CTA, sub-tile and chunk
__host__ __device__ constexpr int gx2_bytes(bool gxh) { return SUB_N * MROWS * (gxh ? 2 : 4); }   // one Gx_t block of a sub-tile: 8 KiB (fp16) / 16 KiB (fp32)
constexpr int CPT2 = SUB_N / 2;                      // accumulator columns per epilogue thread
constexpr int TR_PTS = 32;                           // trace points per (step, sub-tile)
}  // namespace p2

#define AVC_TRACE2(g, pt) do { if (a.trace && blockIdx.x == 0 && blockIdx.y == 0 && t >= TRACE_T0 && t < TRACE_T0 + TRACE_STEPS) \
    a.trace[((t - TRACE_T0) * 2 + (g)) * p2::TR_PTS + (pt)] = clock64(); } while (0)

template <int NKB, bool GXH>
__global__ void __launch_bounds__(p2::THREADS, 1)
ll2_lstm_kernel(const __grid_constant__ CUtensorMap tmW, const LlLstmArgs a) {
    using namespace p2;
    constexpr int NKB_T = NKB < MAX_TMEM_CHUNKS2 ? NKB : MAX_TMEM_CHUNKS2;
    constexpr int NKB_S = NKB - NKB_T;
    constexpr int PSTRIDE = SUB_N * PLANE_ELEMS;          // elements between planes of a 32-row tile of Yx
    static_assert(NKB == 16 || NKB == 8 || NKB == 4, "H must be 1024, 512 or 256");

    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* sW = smem;                                        // NKB_S x 16 KiB
    uint8_t* sH = sW + NKB_S * W_BLK_BYTES;                    // [2 sub-tiles][NKB][2 KiB]
    constexpr int GXB = gx2_bytes(GXH);
    uint8_t* sGx = sH + 2 * NKB * HBLK2;                       // [2 sub-tiles][2 buffers][GXB]: Gx_t is staged ONE STEP AHEAD
    uint8_t* sGT = sGx + 4 * GXB;                              // [2][16 KiB]
    uint64_t* bars = reinterpret_cast<uint64_t*>(sGT + 2 * GT_BYTES);
    uint64_t* wfull = bars;                                    // [1]
    uint64_t* acc_full = bars + 1;                             // [2] the MMAs (both K halves) of a sub-tile's step are complete
    uint64_t* gfull = bars + 4;                                // [2][4] chunk quarter q of sub-tile g is in shared memory (both CTAs)
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 12);

    const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0), lane = threadIdx.x & 31;
    const int ns = blockIdx.x;            // unit slice (32 units)
    const int bt = blockIdx.y;            // pair of 32-utterance sub-tiles
    const uint32_t crank = ptx::cluster_ctarank();
    const size_t tile_elems = (size_t)a.H * SUB_N;                      // one (t, 32-row tile) block of Yx
    const size_t step_elems = tile_elems * a.tiles;

    if (warp == 0) {
        if (lane == 0) {
            if (NKB_S > 0) ptx::prefetch_tmap(&tmW);
            ptx::mbar_init(wfull, 1);
            for (int i = 0; i < 8; ++i) ptx::mbar_init(&gfull[i], 2 * (GROUP_WARPS / 4));   // 2 loader warps per quarter in each CTA
            ptx::mbar_init(&acc_full[0], 2);                                                 // one commit per issuing warp
            ptx::mbar_init(&acc_full[1], 2);
            ptx::fence_barrier_init();
        }
        __syncwarp();
        ptx::tmem_alloc_pair<TMEM_COLS>(tmem_slot);
    }
    ptx::tc_fence_before();
    __syncthreads();
    ptx::cluster_sync();
    ptx::tc_fence_after();
    if (*tmem_slot != 0u) { if (threadIdx.x == 0) printf("autovc_b200: unexpected TMEM base %u\n", *tmem_slot); __trap(); }
    constexpr uint32_t tmem_base = 0u;
    constexpr uint32_t W_COL = D_COLS2;

    // ---- resident W: chunks 0..NKB_T-1 -> tensor memory, the rest -> shared memory ----
    if (warp < 4) {
        const int R = warp * 32 + lane;
        const __nv_bfloat16* wrow = a.Whh + (size_t)(ns * MROWS + R) * a.H;
        const uint32_t wl = tmem_base + W_COL + ((uint32_t)(warp * 32) << 16);
        for (int kc = 0; kc < NKB_T; ++kc) {
            uint4 v[8];
#pragma unroll
            for (int i = 0; i < 8; ++i) v[i] = __ldg(reinterpret_cast<const uint4*>(wrow + kc * KBLK) + i);
            ptx::tmem_st32(wl + (uint32_t)(kc * 32), reinterpret_cast<const uint32_t(&)[32]>(v));
        }
        ptx::tmem_st_wait();
    }
    if (NKB_S > 0 && warp == 4 && lane == 0) {
        ptx::mbar_expect_tx(wfull, (uint32_t)(NKB_S * W_BLK_BYTES));
        for (int i = 0; i < NKB_S; ++i)
            ptx::tma_load_2d(sW + i * W_BLK_BYTES, &tmW, wfull, (NKB_T + i) * KBLK, ns * MROWS);
        ptx::mbar_wait(wfull, 0);
    }
    ptx::tc_fence_before();
    __syncthreads();
    ptx::cluster_sync();
    ptx::tc_fence_after();

    if (warp >= ISSUE_WARP) {
        // ===================== MMA issuers (leader CTA): K half ih of A(t), B(t), A(t+1), ... =====================
        if (crank == 0) {
            const uint32_t sH_u = ptx::smem_u32(sH), sW_u = ptx::smem_u32(sW);
            auto run = [&](auto IH) {
                constexpr int ih = decltype(IH)::value;
                for (int t = 1; t < a.T; ++t) {
                    const uint32_t par = (uint32_t)(t - 1) & 1;
                    auto issue_sub = [&](auto G) {
                        constexpr int g = decltype(G)::value;
                        constexpr uint32_t d_tmem = tmem_base + (uint32_t)((2 * g + ih) * SUB_N);
                        // Issue loop shape: ROLLED over the K chunks with running descriptor words.  ptxas then keeps the whole
                        // sequence on the uniform datapath (UIADD3 / UMOV between the UTCHMMAs); unrolled, it computes every
                        // descriptor in vector registers and pays ~2 R2UR per MMA - 27 cycles per MMA from one warp, against
                        // the 17 the pipe needs at N = 32 (tools/micro/mma_burst.cu, profiles/r2_ll2_lstm_notes.md).
                        auto issue_quarter = [&](auto Q) {
                            constexpr int q = decltype(Q)::value;
                            constexpr int c0 = q * (NKB / 4), c1 = c0 + NKB / 4;
                            constexpr int ts_end = c1 <= NKB_T ? c1 : (c0 < NKB_T ? NKB_T : c0);   // [c0, ts_end) A in TMEM, [ts_end, c1) A in smem
                            ptx::mbar_wait(&gfull[g * 4 + q], par);
                            ptx::tc_fence_after();
                            if (q == 0 && ih == 0 && lane == 0) AVC_TRACE2(g, 8);
                            if (lane == 0) AVC_TRACE2(g, 10 + q);
                            const uint64_t bd0 = ptx::umma_desc_sw128(sH_u + (uint32_t)((g * NKB + c0) * HBLK2));
                            const uint32_t b_hi = (uint32_t)(bd0 >> 32);
                            uint32_t b_lo = (uint32_t)bd0;
                            uint32_t a_t = tmem_base + W_COL + (uint32_t)(c0 * 32);
#pragma unroll 1
                            for (int c = c0; c < ts_end; ++c) {
#pragma unroll
                                for (int k = 0; k < KBLK / 16; ++k) {
                                    ptx::umma2_f16_ts_warp(d_tmem, a_t, ((uint64_t)b_hi << 32) | b_lo, a.idesc, (c != ih * (NKB / 2) || k != 0) ? 1u : 0u);
                                    a_t += 8; b_lo += 2;
                                }
                                b_lo += (uint32_t)(HBLK2 >> 4) - 2 * (KBLK / 16);
                            }
                            if constexpr (ts_end < c1) {
                                const uint64_t ad0 = ptx::umma_desc_sw128(sW_u + (uint32_t)((ts_end - NKB_T) * W_BLK_BYTES));
                                const uint32_t a_hi = (uint32_t)(ad0 >> 32);
                                uint32_t a_lo = (uint32_t)ad0;
#pragma unroll 1
                                for (int c = ts_end; c < c1; ++c) {
#pragma unroll
                                    for (int k = 0; k < KBLK / 16; ++k) {
                                        ptx::umma2_f16_warp(d_tmem, ((uint64_t)a_hi << 32) | a_lo, ((uint64_t)b_hi << 32) | b_lo, a.idesc,
                                                            (c != ih * (NKB / 2) || k != 0) ? 1u : 0u);
                                        a_lo += 2; b_lo += 2;
                                    }
                                    a_lo += (uint32_t)(W_BLK_BYTES >> 4) - 2 * (KBLK / 16);
                                    b_lo += (uint32_t)(HBLK2 >> 4) - 2 * (KBLK / 16);
                                }
                            }
                        };
                        issue_quarter(std::integral_constant<int, 2 * ih>{});
                        issue_quarter(std::integral_constant<int, 2 * ih + 1>{});
                        ptx::umma2_commit_warp(&acc_full[g]);       // -> acc_full[g] of BOTH CTAs
                        if (ih == 1 && lane == 0) AVC_TRACE2(g, 9);
                    };
                    issue_sub(std::integral_constant<int, 0>{});
                    issue_sub(std::integral_constant<int, 1>{});
                }
            };
            if (warp == ISSUE_WARP) run(std::integral_constant<int, 0>{}); else run(std::integral_constant<int, 1>{});
        }
    } else {
        // ===================== sub-tile group g: loaders + epilogue =====================
        const int g = warp >> 3, wg = warp & 7;
        const int tg = (int)(threadIdx.x & (GROUP_THREADS - 1));
        const int b0 = bt * (2 * SUB_N) + g * SUB_N;
        __nv_bfloat16* yx_tile = a.Yx + (size_t)(a.tile0 + bt * 2 + g) * tile_elems;    // + t * step_elems
        uint64_t* my_acc = &acc_full[g];
        // ---- loader geometry: warp wg owns chunks 2wg, 2wg+1; piece i: chunk 2wg + (i>>2), plane 2(i&3) + (lane>>4), row lane&15 ----
        const int lrow = lane & 15, lpl = lane >> 4;
        const uint32_t sH_g = ptx::smem_u32(sH) + (uint32_t)(g * NKB * HBLK2);
        constexpr int CPW = NKB / GROUP_WARPS > 0 ? NKB / GROUP_WARPS : 1;      // chunks per warp: 2 (H=1024), 1 (H=512)
        constexpr int WPC2 = NKB >= GROUP_WARPS ? 1 : GROUP_WARPS / NKB;          // warps per chunk: 1, or 2 (H=256)
        constexpr int NP = SUB_ROWS * 8 * NKB / GROUP_THREADS;                  // 16-byte pieces per lane and step: 8 / 4 / 2
        auto piece_chunk = [&](int i) { return WPC2 == 1 ? wg * CPW + (i >> 2) : wg / WPC2; };
        auto piece_plane = [&](int i) { return WPC2 == 1 ? 2 * (i & 3) + lpl : (wg % WPC2) * (8 / WPC2) + 2 * i + lpl; };
        const __nv_bfloat16* rows_base = yx_tile + (size_t)(crank * SUB_ROWS + lrow) * PLANE_ELEMS;
        auto piece_src = [&](int i) { return rows_base + (size_t)(piece_chunk(i) * 8 + piece_plane(i)) * PSTRIDE; };
        const int gq = (wg * 4) / GROUP_WARPS;                  // chunk quarter this warp's chunks belong to (warps 2q, 2q+1)

        // ---- epilogue geometry ----
        const int q = wg & 3, half = wg >> 2;                 // TMEM lane quarter (= warp % 4); column half (16 utterances)
        const int R = q * 32 + lane;                          // gate row of this CTA = TMEM lane
        const int gj = lane & 3;                              // gate of this lane: 0=i 1=f 2=g 3=o
        const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(2 * g * SUB_N + half * CPT2);   // K half 1: + SUB_N
        const float act_s = gj == 2 ? 2.0f : 1.0f, act_o = gj == 2 ? -1.0f : 0.0f;
        const uint32_t gt_smem = ptx::smem_u32(sGT) + (uint32_t)(g * GT_BYTES + wg * (16 * 32 * 4));   // this warp's transpose tile
        float c[CPT2 / 4];
#pragma unroll
        for (int k = 0; k < CPT2 / 4; ++k) c[k] = 0.0f;
        const uint32_t gx_smem0 = ptx::smem_u32(sGx) + (uint32_t)(g * 2 * GXB);
        // Gx_t staging, one step ahead: the block of step t is copied into buffer t & 1 while step t-1 runs (the copies come from
        // HBM: issued in their own step they would not be back before the MMAs of a 32-wide sub-tile are).  WARP-PRIVATE: a warp
        // copies exactly the [16 utterances of its half] x [32 gate rows of its lane quarter] block its own lanes read, so the
        // hand-over needs a __syncwarp, not a barrier of the group.  One commit group per step (an empty one past the end), so
        // "all but the newest group" is always "this step's block".  The row index t / gx_div is kept incrementally.
        int gx_row = 0, gx_left = a.gx_div;       // row of the NEXT block to stage, steps left on that row
        auto stage_gx = [&](int t) {
            if (t < a.T) {
                const size_t toff = (size_t)gx_row * a.gx_st + (size_t)ns * MROWS + (size_t)(q * 32);
                constexpr int ES = GXH ? 2 : 4, SEGS = 32 * ES / 16;          // 16-byte pieces per utterance of the warp's block: 4 / 8
                const uint32_t dst = gx_smem0 + (uint32_t)((t & 1) * GXB);
#pragma unroll
                for (int i = 0; i < CPT2 * SEGS / 32; ++i) {
                    const int u = half * CPT2 + lane / SEGS + (32 / SEGS) * i, seg = lane % SEGS;
                    const int b = b0 + u < a.B ? b0 + u : a.B - 1;
                    cp_async16(dst + (uint32_t)((u * MROWS + q * 32) * ES + seg * 16),
                               reinterpret_cast<const char*>(a.Gx) + ((size_t)b * a.gx_sb + toff) * ES + seg * 16);
                }
                if (--gx_left == 0) { ++gx_row; gx_left = a.gx_div; }
            }
            cp_async_commit();
        };
        uint32_t gx_smem = gx_smem0;
        auto gx_at = [&](int k) -> float {
            if constexpr (GXH) {
                uint16_t hbits;
                asm volatile("ld.shared.u16 %0, [%1];" : "=h"(hbits) : "r"(gx_smem + (uint32_t)(((half * CPT2 + k) * MROWS + R) * 2)));
                return __half2float(__ushort_as_half(hbits));
            } else {
                return ld_shared_f32(gx_smem + (uint32_t)(((half * CPT2 + k) * MROWS + R) * 4));
            }
        };
        stage_gx(0);
        stage_gx(1);

        for (int t = 0; t < a.T; ++t) {
            gx_smem = gx_smem0 + (uint32_t)((t & 1) * GXB);
            if (t >= 1) {
                // ---- h_{t-1} of this sub-tile: L2 -> registers (poll) -> swizzled operand tile ----
                const size_t soff = (size_t)(t - 1) * step_elems;
                if (tg == 0) AVC_TRACE2(g, 0);
                uint4 v[NP];
#pragma unroll
                for (int i = 0; i < NP; ++i) v[i] = ld_strong(piece_src(i) + soff);
                for (uint32_t spin = 0;; ++spin) {          // pieces not written yet: fetch all of them again, together
                    bool missing = false;
#pragma unroll
                    for (int i = 0; i < NP; ++i)
                        if (unwritten(v[i])) { v[i] = ld_strong(piece_src(i) + soff); missing = true; }
                    if (!missing) break;
                    if (spin > (1u << 22)) {
                        printf("autovc_b200: lstm h poll timed out (slice %d tile %d sub %d t %d warp %d lane %d)\n", ns, bt, g, t, wg, lane);
                        __trap();
                    }
                }
                if (tg == 0) AVC_TRACE2(g, 1);
#pragma unroll
                for (int i = 0; i < NP; ++i) {
                    const int pl = piece_plane(i);
                    st_shared_v4(sH_g + (uint32_t)(piece_chunk(i) * HBLK2 + lrow * 128 + ((pl ^ (lrow & 7)) << 4)), v[i]);
                }
                ptx::fence_proxy_async_smem();
                __syncwarp();
                if (lane == 0) ptx::mbar_arrive_cluster(&gfull[g * 4 + gq], 0u);     // the LEADER's barrier (both CTAs arrive)
                if (tg == 0) AVC_TRACE2(g, 2);
                if (lane == 0) AVC_TRACE2(g, 16 + wg);
                stage_gx(t + 1);
            }
            // ---- epilogue: gates, cell state, h_t ----
            if (tg == 0) AVC_TRACE2(g, 24);
            asm volatile("cp.async.wait_group 1;" ::: "memory");     // own pieces of Gx_t (the newest group is Gx_{t+1}) ...
            __syncwarp();                                            // ... and the other lanes' (the block is this warp's alone)
            if (tg == 0) AVC_TRACE2(g, 26);
            float act[CPT2];
            if (t >= 1) {
                ptx::mbar_wait(my_acc, (uint32_t)(t - 1) & 1);
                if (tg == 0) AVC_TRACE2(g, 3);
                ptx::tc_fence_after();
                uint32_t acc0[16], acc1[16];
                ptx::tmem_ld16(taddr, acc0);
                ptx::tmem_ld16(taddr + (uint32_t)SUB_N, acc1);
                ptx::tmem_ld_wait();
                ptx::tc_fence_before();
                if (tg == 0) AVC_TRACE2(g, 5);
#pragma unroll
                for (int k = 0; k < CPT2; ++k) act[k] = __uint_as_float(acc0[k]) + __uint_as_float(acc1[k]);
            } else {
#pragma unroll
                for (int k = 0; k < CPT2; ++k) act[k] = 0.0f;
            }
            // gate activation of this lane's row (x' already is the tanh argument, see pack_lstm), written to the warp's transpose
            // tile [column][gate row]: the four gates of a unit are four consecutive floats there
            if (a.approx) {
                const float ps = gj == 2 ? 1.0f : 0.5f, po = gj == 2 ? 0.0f : 0.5f;
#pragma unroll
                for (int k = 0; k < CPT2; ++k) {
                    const float x = act[k] + gx_at(k);
                    float th;
                    asm("tanh.approx.f32 %0, %1;" : "=f"(th) : "f"(x));
                    act[k] = fmaf(ps, th, po);
                }
            } else {
#pragma unroll
                for (int k = 0; k < CPT2; ++k) {
                    const float x = act[k] + gx_at(k);
                    act[k] = __fdividef(act_s, 1.0f + __expf(-2.0f * x)) + act_o;
                }
            }
#pragma unroll
            for (int k = 0; k < CPT2; ++k)      // after ALL the loads above: a store in the loop would order every Gx read behind it
                asm volatile("st.shared.f32 [%0], %1;" ::"r"(gt_smem + (uint32_t)((k * 32 + lane) * 4)), "f"(act[k]) : "memory");
            __syncwarp();
            if (tg == 0) AVC_TRACE2(g, 6);
            // lane j: unit (j >> 2) of this quarter, columns 4 (j & 3) .. + 3 of this half: i, f, g, o as one 16-byte read each
            constexpr int H2 = CPT2 / 4;
            uint32_t hp[H2 / 2];
#pragma unroll
            for (int k = 0; k < H2; k += 2) {
                float hv[2];
#pragma unroll
                for (int w = 0; w < 2; ++w) {
                    const int kk = k + w;
                    const uint4 gates = ld_shared_v4(gt_smem + (uint32_t)(((gj * H2 + kk) * 32 + (lane & ~3)) * 4));
                    const float ig = __uint_as_float(gates.x), fg = __uint_as_float(gates.y), gg = __uint_as_float(gates.z), og = __uint_as_float(gates.w);
                    c[kk] = cell_update(fg, c[kk], ig, gg);
                    float tc;
                    if (a.approx) asm("tanh.approx.f32 %0, %1;" : "=f"(tc) : "f"(c[kk]));
                    else tc = 1.0f - __fdividef(2.0f, 1.0f + __expf(2.0f * c[kk]));
                    hv[w] = og * tc;
                }
                __nv_bfloat162 h2 = __floats2bfloat162_rn(hv[0], hv[1]);
                hp[k >> 1] = *reinterpret_cast<uint32_t*>(&h2);
            }
            __syncwarp();        // the tile is rewritten by the next step's activations
            if (tg == 0) AVC_TRACE2(g, 7);
            // h_t out, straight from registers.  The lane holds one unit x four utterances (hp[0] = columns 0,1; hp[1] = columns 2,3
            // of its block); lanes j and j ^ 4 hold neighbouring units of the same utterances.  One exchange gives every lane a
            // (unit pair) x (two utterances) block, i.e. two 32-bit words: the validity protocol needs whole 32-bit words from one
            // thread (a word with one stale half would not read as "unwritten").  No shared-memory staging, no barrier.
            {
                const bool odd = (lane >> 2) & 1;
                const uint32_t mine = odd ? hp[1] : hp[0];
                const uint32_t theirs = __shfl_xor_sync(0xffffffffu, odd ? hp[0] : hp[1], 4);
                const uint32_t ev = odd ? theirs : mine, od = odd ? mine : theirs;      // even unit / odd unit, same two utterances
                const uint32_t w0 = __byte_perm(ev, od, 0x5410), w1 = __byte_perm(ev, od, 0x7632);
                const int col = half * CPT2 + gj * H2 + (odd ? 2 : 0);                  // first of the two utterances (within the sub-tile)
                const int upair = (lane >> 3) * 2;                                       // even unit of the pair, within this quarter's 8
                __nv_bfloat16* yx = yx_tile + (size_t)t * step_elems + ((size_t)(ns * 4 + q) * SUB_N + col) * PLANE_ELEMS + upair;
                if (t + 1 < a.T) {
                    asm volatile("st.relaxed.gpu.global.u32 [%0], %1;" ::"l"(yx), "r"(w0) : "memory");
                    asm volatile("st.relaxed.gpu.global.u32 [%0], %1;" ::"l"(yx + PLANE_ELEMS), "r"(w1) : "memory");
                }
                __nv_bfloat16* y = a.Y + (size_t)t * a.y_st + (size_t)ns * SLICE_UNITS + q * 8 + upair;
                if (b0 + col < a.B) *reinterpret_cast<uint32_t*>(y + (size_t)(b0 + col) * a.y_sb) = w0;
                if (b0 + col + 1 < a.B) *reinterpret_cast<uint32_t*>(y + (size_t)(b0 + col + 1) * a.y_sb) = w1;
            }
            if (tg == 0) AVC_TRACE2(g, 4);
        }
    }

    ptx::tc_fence_before();
    __syncthreads();
    ptx::cluster_sync();
    if (warp == 0) {
        ptx::tc_fence_after();
        ptx::tmem_dealloc_pair<TMEM_COLS>(tmem_base);
    }
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
EncodeTiledFn g_encode_ll = nullptr;

template <int TN>
const void* kernel_for_tn(int nkb, bool gxh) {
    switch (nkb) {
        case 16: return gxh ? (const void*)ll_lstm_kernel<16, true, TN> : (const void*)ll_lstm_kernel<16, false, TN>;
        case 8: return gxh ? (const void*)ll_lstm_kernel<8, true, TN> : (const void*)ll_lstm_kernel<8, false, TN>;
        default: return gxh ? (const void*)ll_lstm_kernel<4, true, TN> : (const void*)ll_lstm_kernel<4, false, TN>;
    }
}
const void* kernel_for(int nkb, bool gxh, int tn) { return tn == 32 ? kernel_for_tn<32>(nkb, gxh) : kernel_for_tn<64>(nkb, gxh); }

const void* kernel_for2(int nkb, bool gxh) {
    switch (nkb) {
        case 16: return gxh ? (const void*)ll2_lstm_kernel<16, true> : (const void*)ll2_lstm_kernel<16, false>;
        case 8: return gxh ? (const void*)ll2_lstm_kernel<8, true> : (const void*)ll2_lstm_kernel<8, false>;
        default: return gxh ? (const void*)ll2_lstm_kernel<4, true> : (const void*)ll2_lstm_kernel<4, false>;
    }
}
int smem_for2(int nkb, bool gxh) {
    const int nkb_s = nkb > p2::MAX_TMEM_CHUNKS2 ? nkb - p2::MAX_TMEM_CHUNKS2 : 0;
    return nkb_s * W_BLK_BYTES + 2 * nkb * p2::HBLK2 + 4 * p2::gx2_bytes(gxh) + 2 * p2::GT_BYTES + 1024 + 512;
}

int smem_for(int nkb) {
    const int nkb_s = nkb > MAX_TMEM_CHUNKS ? nkb - MAX_TMEM_CHUNKS : 0;
    return nkb_s * W_BLK_BYTES + nkb * H_BLK_BYTES + GX_BYTES + HT_BYTES + 1024 + 512;
}

}  // namespace

struct LlLstmPlan {
    CUtensorMap tmW;
    std::vector<LlLstmArgs> args;
    std::vector<dim3> grids;
    int nkb, tn;
    bool pipelined;           // ll2_lstm_kernel: two 32-utterance sub-tiles per CTA pair
    bool gx_f16;
    int smem_bytes;
    void* yx_base; size_t yx_bytes;       // region to pre-fill with the "unwritten" pattern before every launch
};

const char* ll_lstm_init() {
    if (g_encode_ll) return nullptr;
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult qres;
    cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres);
    if (e != cudaSuccess || qres != cudaDriverEntryPointSuccess || !fn) return "cuTensorMapEncodeTiled not available from the driver";
    for (int i = 0; i < 12; ++i)
        if (cudaFuncSetAttribute(kernel_for(4 << ((i >> 1) % 3), i & 1, i < 6 ? 64 : 32), cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 smem_for(4 << ((i >> 1) % 3))) != cudaSuccess)
            return "cudaFuncSetAttribute(MaxDynamicSharedMemorySize) failed for ll_lstm_kernel";
    for (int i = 0; i < 6; ++i) {
        const int bytes = smem_for2(4 << (i >> 1), i & 1);
        if (bytes > 227 * 1024) continue;          // H = 1024 with fp32 Gx does not fit: that plan takes the one-tile kernel
        if (cudaFuncSetAttribute(kernel_for2(4 << (i >> 1), i & 1), cudaFuncAttributeMaxDynamicSharedMemorySize, bytes) != cudaSuccess)
            return "cudaFuncSetAttribute(MaxDynamicSharedMemorySize) failed for ll2_lstm_kernel";
    }
    g_encode_ll = reinterpret_cast<EncodeTiledFn>(fn);
    return nullptr;
}

size_t ll_lstm_exchange_bytes(int B, int T, int H) { return (size_t)T * ceil_div(B, TILE_N) * TILE_N * H * 2; }

LlLstmPlan* ll_lstm_plan_create(const void* Whh_bf16, const void* Gx, int gx_type, int Tg, int gx_div, void* Y, void* Yx, int B, int T, int H,
                                int num_sms, const char** err) {
    static thread_local char msg[256];
    auto bad = [&](const char* m) -> LlLstmPlan* { if (err) *err = m; return nullptr; };
    if (!g_encode_ll) return bad("ll_lstm_init() not called");
    if (H != 1024 && H != 512 && H != 256) return bad("hidden size must be 256, 512 or 1024");
    if (gx_type != DT_F32 && gx_type != DT_F16) return bad("Gx must be fp32 or fp16");
    const int G_n = H / SLICE_UNITS, nkb = H / KBLK;
    if (G_n > num_sms) return bad("hidden size too large for one co-resident grid");
    LlLstmPlan* plan = new LlLstmPlan();
    plan->nkb = nkb;
    plan->smem_bytes = smem_for(nkb);
    plan->gx_f16 = gx_type == DT_F16;
    // tile width: 32 utterances per CTA pair when twice the tiles still fit the machine (small batches: shorter steps)
    int tn = ceil_div(B, 32) * G_n <= (num_sms & ~1) ? 32 : 64;
    if (const char* tv = getenv("AVC_LSTM_TN")) tn = atoi(tv) == 32 ? 32 : 64;
    // Batches that would take 64-wide tiles run as two pipelined 32-wide sub-tiles per CTA pair instead (ll2_lstm_kernel): one
    // sub-tile's exchange latency hides behind the other's MMAs and epilogue.  AVC_LSTM_P2=0/1 forces the choice (A/B aid).
    bool pipelined = tn == 64;
    if (const char* pv = getenv("AVC_LSTM_P2")) pipelined = atoi(pv) != 0;
    if (pipelined && smem_for2(nkb, plan->gx_f16) > 227 * 1024) pipelined = false;     // H = 1024 with fp32 Gx (AVC_GX_F32): no room for the second Gx buffer
    plan->pipelined = pipelined;
    if (pipelined) { tn = 64; plan->smem_bytes = smem_for2(nkb, plan->gx_f16); }
    plan->tn = tn;
    plan->yx_base = Yx; plan->yx_bytes = (size_t)T * ceil_div(B, tn) * tn * H * 2;
    {
        cuuint64_t gdim[2] = {(cuuint64_t)H, (cuuint64_t)4 * H};
        cuuint64_t gstr[1] = {(cuuint64_t)H * 2};
        cuuint32_t box[2] = {(cuuint32_t)KBLK, (cuuint32_t)MROWS};
        cuuint32_t est[2] = {1, 1};
        CUresult r = g_encode_ll(&plan->tmW, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(Whh_bf16), gdim, gstr, box, est,
                                 CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                                 CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (r != CUDA_SUCCESS) { delete plan; snprintf(msg, sizeof msg, "cuTensorMapEncodeTiled(Whh) failed: %d", (int)r); return bad(msg); }
    }
    const int tiles_per_launch = (num_sms & ~1) / G_n;
    const int rows_per_launch = tiles_per_launch * tn;
    for (int bs = 0; bs < B; bs += rows_per_launch) {
        const int Bc = (B - bs) < rows_per_launch ? (B - bs) : rows_per_launch;
        LlLstmArgs a{};
        a.B = Bc; a.T = T; a.H = H; a.gx_div = gx_div;
        a.y_sb = (long long)T * H; a.y_st = H; a.gx_sb = (long long)Tg * 4 * H; a.gx_st = 4LL * H;
        a.Gx = (const char*)Gx + (size_t)bs * a.gx_sb * dtype_size(gx_type);
        a.Whh = reinterpret_cast<const __nv_bfloat16*>(Whh_bf16);
        a.Y = reinterpret_cast<__nv_bfloat16*>(Y) + (size_t)bs * a.y_sb;
        a.Yx = reinterpret_cast<__nv_bfloat16*>(Yx);
        a.tiles = ceil_div(B, tn); a.tile0 = bs / tn;
        if (pipelined) { a.tiles *= 2; a.tile0 *= 2; }           // Yx is indexed in 32-row tiles
        a.approx = getenv("AVC_LSTM_EXACT") == nullptr;
        // canary poll of the loader-only warps: worth 10 % at H = 1024 with a full grid (without it they hammer the fabric with
        // 64 KiB waves per CTA while the epilogue runs); at H <= 512 (16 KiB waves) and on small grids it only adds a round trip
        // per step (lstm1 at B = 256: 0.328 -> 0.308 ms; at B <= 32: 0.313 -> 0.287 ms)
        { const char* cv = getenv("AVC_LSTM_CANARY"); a.canary = cv ? atoi(cv) : (H <= 512 || G_n * ceil_div(Bc, tn) <= 16 ? 0 : 1); }
        { const char* gv = getenv("AVC_LSTM_GX_LATE"); a.gx_late = gv ? atoi(gv) : 1; }
        a.idesc = ptx::umma_idesc_f16(1, 2 * MROWS, pipelined ? p2::SUB_N : tn);   // the pair's MMA: M = 256 (128 gate rows per CTA), N = tile width
        plan->args.push_back(a);
        plan->grids.push_back(dim3(G_n, ceil_div(Bc, tn)));
    }
    return plan;
}

void ll_lstm_plan_destroy(LlLstmPlan* plan) { delete plan; }

cudaError_t launch_ll_lstm(const LlLstmPlan* plan, cudaStream_t s, int* launches) {
    static long long* d_trace = nullptr;
    const bool tracing = getenv("AVC_LSTM_TRACE") != nullptr;
    if (tracing && !d_trace) cudaMalloc(&d_trace, sizeof(long long) * TRACE_WORDS);
    // "unwritten" pattern: every 16-bit element = 0xFFFF
    cudaError_t e = cudaMemsetAsync(plan->yx_base, 0xFF, plan->yx_bytes, s);
    if (e != cudaSuccess) return e;
    for (size_t i = 0; i < plan->args.size(); ++i) {
        LlLstmArgs args = plan->args[i];
        args.trace = tracing ? d_trace : nullptr;
        if (tracing) cudaMemsetAsync(d_trace, 0, sizeof(long long) * TRACE_WORDS, s);
        void* kargs[2] = {(void*)&plan->tmW, (void*)&args};
        // CTAs wait on one another's output: cooperative launch (guaranteed co-residency), with the checked non-cooperative
        // retry of launch_persistent() for profilers (AVC_LSTM_NOCOOP forces it)
        e = launch_persistent(plan->pipelined ? kernel_for2(plan->nkb, plan->gx_f16) : kernel_for(plan->nkb, plan->gx_f16, plan->tn), plan->grids[i],
                              dim3(plan->pipelined ? p2::THREADS : NUM_THREADS), (size_t)plan->smem_bytes, 2, kargs, s,
                              getenv("AVC_LSTM_NOCOOP") != nullptr);
        if (e != cudaSuccess) return e;
        if (launches) ++*launches;
        if (tracing && plan->pipelined) {
            long long hbuf[TRACE_WORDS];
            cudaStreamSynchronize(s);
            cudaMemcpy(hbuf, d_trace, sizeof hbuf, cudaMemcpyDeviceToHost);
            fprintf(stderr, "ll2_lstm trace H=%d B=%d grid=(%d,%d): per sub-tile, cycles since its loads were issued (warp 0 of the group unless noted): landed published | "
                    "quarters ready q0..q3 (issuers), committed | Gx wait entered, Gx ready | acc_seen tmem_read gates cell stored | publish time of warps 0..7 || absolute start\n",
                    args.H, args.B, plan->grids[i].x, plan->grids[i].y);
            const long long t00 = hbuf[0];
            for (int st = 0; st < TRACE_STEPS; ++st)
                for (int g = 0; g < 2; ++g) {
                    const long long* r = hbuf + (st * 2 + g) * p2::TR_PTS;
                    if (!r[0]) continue;
                    const long long z = r[0];
                    fprintf(stderr, "  t=%d sub %d: %5lld %5lld | %5lld %5lld %5lld %5lld, %5lld | %5lld %5lld | %5lld %5lld %5lld %5lld %5lld | %5lld %5lld %5lld %5lld %5lld %5lld %5lld %5lld || %7lld\n",
                            TRACE_T0 + st, g, r[1] - z, r[2] - z, r[10] - z, r[11] - z, r[12] - z, r[13] - z, r[9] - z, r[24] - z, r[26] - z, r[3] - z, r[5] - z, r[6] - z, r[7] - z, r[4] - z,
                            r[16] - z, r[17] - z, r[18] - z, r[19] - z, r[20] - z, r[21] - z, r[22] - z, r[23] - z, z - t00);
                }
        } else if (tracing) {   // debugging aid: synchronous dump of the per-step timeline of CTA (0,0)
            long long hbuf[TRACE_WORDS];
            cudaStreamSynchronize(s);
            cudaMemcpy(hbuf, d_trace, sizeof hbuf, cudaMemcpyDeviceToHost);
            fprintf(stderr, "ll_lstm trace H=%d B=%d grid=(%d,%d): cycles since warp 0 saw its canary: "
                    "warp0_done mma_issue_start mma_issued acc_seen h_stored (activations cell_update staged) | step period\n",
                    args.H, args.B, plan->grids[i].x, plan->grids[i].y);
            for (int st = 0; st < TRACE_STEPS; ++st) {
                const long long* r = hbuf + st * TRACE_PTS;
                if (!r[0]) continue;
                fprintf(stderr, "  t=%d:", TRACE_T0 + st);
                for (int p = 1; p <= 5; ++p) fprintf(stderr, " %6lld", r[p] - r[0]);
                fprintf(stderr, " (%lld %lld %lld)", r[6] - r[0], r[7] - r[0], r[8] - r[0]);
                fprintf(stderr, " [landed %lld groups %lld %lld %lld %lld issued0 %lld]", r[9] - r[0], r[10] - r[0], r[11] - r[0], r[12] - r[0], r[13] - r[0], r[14] - r[0]);
                if (st + 1 < TRACE_STEPS && r[TRACE_PTS]) fprintf(stderr, " | %6lld", r[TRACE_PTS] - r[0]);
                fprintf(stderr, "\n");
            }
        }
    }
    return cudaSuccess;
}

}  // namespace avc
