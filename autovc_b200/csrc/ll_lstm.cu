// Persistent weight-stationary LSTM recurrence on the tensor cores (sm_100a): the decoder's lstm1 (H=512) and lstm2 (H=1024, two
// layers run one after the other) [reference decoder.py:37,53,65,74] and the speaker encoder's three layers (H=256).
// The input projection x_t W_ih^T + b_ih + b_hh for all t is hoisted into one dense GEMM (tc_gemm.cu) beforehand.
//
// Decomposition: grid = (H/32 unit slices) x (batch tiles), all CTAs co-resident (cooperative launch, one CTA per SM), launched
// as clusters of 2 CTAs = two adjacent unit slices of one batch tile.  A step's GEMM is computed TRANSPOSED,
//   D[128 gate rows x 32 utterances] (TMEM, fp32) = W_slice[128 x H] . h_{t-1}^T[H x 32]     (cta_group::2: M = 256 over the pair)
// so that the resident operand is A and the streamed operand is B:
//   * A = the CTA's 128 gate rows (32 hidden units x i,f,g,o, unit-major packing) of W_hh, resident for the whole sequence: in
//     TENSOR MEMORY as far as it fits beside the accumulators (12 or 14 of the 16 K chunks at H = 1024), the rest in shared
//     memory.  An MMA with A in tensor memory reads only a 1 KiB h block from shared memory and retires in ~17 cycles at N = 32
//     (40 with A in shared memory; tools/micro/mma_rate.cu).
//   * B = h_{t-1} of the tile's 32 utterances, split over the pair's shared memory (16 rows each): a CTA ingests 32 KiB per tile
//     and step at H = 1024.
//
// A step of one tile is a dependent chain  exchange (h_t stored -> visible in L2 -> load wave)  ->  MMAs  ->  gate epilogue  in
// which every unit idles while the others work, and with all 128 gate-row slices of 256 utterances spread over the machine the
// chain's LATENCY, not any unit's throughput, is the step time.  Two things follow.
//   (1) NSUB = 2 (batches that need 64 utterances per CTA pair to fit the machine, e.g. lstm2 at B = 256): TWO independent
//       32-utterance sub-tiles per pair over the same resident W_hh; while sub-tile A's h crosses L2, sub-tile B's MMAs and
//       epilogue run.  NSUB = 1 otherwise (every tile gets its own pairs; nothing to pipeline with).
//   (2) No CTA-level barrier inside the time loop: a warp waits for its own h pieces, for the accumulator mbarrier, and for
//       nothing else.  (With group barriers 15 % of all warp samples sat behind the group's slowest poller and 5 % in front of
//       the stores; profiles/r2_ll_lstm_notes.md.)  Hence the Gx_t block is staged warp-privately, the 4x4 gate transpose goes
//       through a per-warp shared-memory tile, and h_t leaves straight from registers.
//   (3) The chain is serial code on warps that share a scheduler four or five ways: every instruction on a warp's path through a
//       step costs ~9 cycles of the step (ncu source counters; profiles/r2_ll_lstm_notes.md).  Hence: nothing that does not
//       depend on t is computed inside the time loop (ring / output / Gx source pointers, the remote barrier address and the
//       accumulator address are built once and kept opaque to the compiler), Gx_t is fetched before the accumulator wait,
//       the epilogue arithmetic is packed fp32x2, the transpose tile is bank-conflict free, and tracing has its own instantiation.
//
// Warp roles (NSUB x GW + 2 warps; GW = 8 group warps per sub-tile, 16 for the one sub-tile of an H = 1024 CTA - each warp then
// owns half the chunks and half the columns):
//   warps GW g..GW g + GW - 1   sub-tile g: loaders of its h_{t-1} rows (poll -> swizzled K-major operand tile -> fence.proxy.async ->
//                    arrive on the mbarrier of its own K part in the leader CTA) AND its gate epilogue (tcgen05.ld of 16 or 8 columns per
//                    thread, + Gx_t, MUFU.TANH gates, cell update with c in registers, h_t stores)
//   last two warps   MMA issuers (leader CTA), one K half each into its own accumulator, each strictly alternating
//                    A(t), B(t), A(t+1), ... in program order.  That order (a) keeps the accumulation order independent of timing
//                    (results are bitwise reproducible, and a row's result does not depend on NSUB or on the batch around it)
//                    and (b) pins the two chains one MMA phase apart: two free-running chains fall into lockstep and then share
//                    every unit at half rate, which is one 64-wide tile again (profiles/r1_ll_lstm_step_timeline.md).
//                    Issue loops are ROLLED over the K chunks with running descriptor words: ptxas then keeps the sequence on
//                    the uniform datapath; unrolled, it builds every descriptor in vector registers and pays ~2 R2UR per MMA.
//
// Exchange protocol ("the data is the flag"): CTAs exchange h through an L2-resident RING of two step slots,
// Yx[2][32-utterance tile][H/8 planes][32 rows][8 bf16] (1 MiB at B = 256, H = 1024).  |h| <= 1 (h = o tanh(c)), so bit 14 of a
// bf16 h - the top exponent bit, set only for |x| >= 2, Inf and NaN - is always 0 and carries a one-bit EPOCH instead: step t
// writes slot t & 1 with epoch (t >> 1) & 1.  A slot's previous content (step t-2) has the opposite epoch, the pre-fill (0xFF,
// once per launch, 1 MiB) reads as epoch 1 for the epoch-0 steps 0 and 1: "not written yet" is "wrong epoch in any 32-bit word"
// (32-bit accesses are single-copy atomic and both halves of a word come from one thread).  Producers store h_t as tagged
// 32-bit words; consumers read with strong (L1-bypassing) coalesced 16-byte loads, fetch the missing pieces again, together,
// and strip the tag on the way into the operand tile.  No step counter, no release/acquire fence pair, no second round trip
// between "flag seen" and "data fetched", and no T-sized buffer to pre-fill and write back (67 MB per lstm2 layer before).
// Two slots suffice: a CTA that stores h_t has consumed h_{t-1} from every producer of its tile, so every one of them had finished
// step t-1 and therefore had read all of h_{t-2} - the content the store replaces; and no CTA can be two steps ahead of a reader.
// A non-finite h (NaN / Inf in the inputs) would carry bit 14 itself: the exchange copy then gets +-1 in its place (the
// recurrence stays alive) while the output copy Y keeps the true value, so the NaN still reaches the results.
// No acc_empty barrier either: a sub-tile's operands for step t+1 are published by the very warps that read its accumulators of
// step t, so "operands complete" implies "accumulators drained" in both CTAs.
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
constexpr int MROWS = 4 * SLICE_UNITS;    // gate rows per CTA = UMMA M per CTA = TMEM lanes
constexpr int KBLK = 64;                  // K elements per chunk (128 bytes per operand row)
constexpr int W_BLK_BYTES = MROWS * KBLK * 2;     // 16 KiB (shared-memory part of W)
constexpr int PLANE_ELEMS = 8;            // Yx plane width: 8 bf16 = 16 B
constexpr uint32_t TMEM_COLS = 512;
constexpr uint32_t EPOCH_BITS = 0x40004000u;          // bit 14 of both bf16 halves of a 32-bit word: the step epoch of the exchange ring
constexpr int SUB_N = 32;                 // utterances per (sub-)tile = UMMA N
constexpr int SUB_ROWS = SUB_N / 2;       // rows of h staged per CTA and sub-tile
// loader + epilogue warps per sub-tile: 8, or 16 when an H = 1024 CTA has ONE sub-tile (half the columns, pieces and
// transcendental chain per warp on the step's critical path: config 3 lstm2 4.45 -> 4.28 ms, B = 1 p50 -1.5 %; at H = 512 the
// extra pollers cost more than they return: 1.755 -> 1.770 ms)
__host__ __device__ constexpr int group_warps(int nsub, int nkb) { return nsub == 1 && nkb == 16 ? 16 : 8; }
constexpr int GT_BYTES = SUB_N * MROWS * 4;          // 16 KiB per sub-tile: the group's [32 columns][128 gate rows] fp32 transpose tiles
constexpr int HBLK2 = SUB_ROWS * KBLK * 2;           // 2 KiB per CTA, sub-tile and chunk of the operand tile
__host__ __device__ constexpr int threads(int nsub, int nkb) { return (nsub * group_warps(nsub, nkb) + 2) * 32; }    // + the two issuer warps: 320 / 576
__host__ __device__ constexpr int max_tmem_chunks(int nsub) { return (512 - 2 * nsub * SUB_N) / 32; }   // two accumulators per sub-tile: 14 / 12
__host__ __device__ constexpr int gx2_bytes(bool gxh) { return SUB_N * MROWS * (gxh ? 2 : 4); }         // one Gx_t block of a sub-tile: 8 / 16 KiB

struct LlLstmArgs {
    int B, T, H;
    int gx_div;                  // Gx row for frame t is t / gx_div (lstm1 reads the de-duplicated projection)
    const void* Gx;              // [B][Tg][4H] fp32 (or fp16: GXH kernels), unit-major gate order, bias included
    const __nv_bfloat16* Whh;    // [4H][H] bf16, unit-major gate rows (sigmoid rows pre-scaled by 0.5, api.cu::pack_lstm)
    __nv_bfloat16* Y;            // [B][T][H] bf16 output sequence for the next layer
    __nv_bfloat16* Yx;           // exchange ring: [2][tiles][H/8][32][8] bf16 (epoch-tagged), pre-filled with 0xFF
    int tiles, tile0;            // 32-utterance tiles in Yx, first tile of this launch
    long long y_sb, y_st;        // element strides of Y along utterance / frame
    long long gx_sb, gx_st;      // same for Gx rows (frame index t / gx_div)
    uint32_t idesc;              // M = 256, N = 32
    uint32_t idesc16;            // M = 256, N = 16: sub-tiles with at most 8 utterances (rows 0..7 of the leader CTA's operand tile)
    int approx;                  // 1 (default): gate activations with MUFU.TANH (tanh.approx.f32, abs err ~5e-4: below the bf16 rounding of h;
                                 // measured end-to-end error unchanged); AVC_LSTM_EXACT=1 selects ex2/rcp based sigmoid/tanh
    long long* trace;            // optional clock64 trace of CTA (0,0) (AVC_LSTM_TRACE=1), nullptr normally
};

constexpr int TRACE_T0 = 8, TRACE_STEPS = 8, TR_PTS = 32;        // optional clock64 trace: steps 8..15 of CTA (0,0), 32 stamps per (step, sub-tile)
constexpr int TRACE_WORDS = (2 * TRACE_STEPS + 1) * TR_PTS;
__device__ __forceinline__ uint4 ld_strong(const void* p) {
    uint4 v;
    asm volatile("ld.relaxed.gpu.global.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ bool unwritten(const uint4& v, uint32_t etag) {      // etag: EPOCH_BITS or 0, the epoch the piece must carry
    return (((v.x ^ etag) | (v.y ^ etag) | (v.z ^ etag) | (v.w ^ etag)) & EPOCH_BITS) != 0u;
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
__device__ __forceinline__ void cp_async16(uint32_t dst_smem, const void* src) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst_smem), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
// c' = f c + i g with ONE fixed rounding order in every kernel of this file (rows of a batch must not depend on which kernel
// variant the batch size selects): the product i g is rounded, then fused into f c
__device__ __forceinline__ float cell_update(float fg, float c, float ig, float gg) { return fmaf(fg, c, __fmul_rn(ig, gg)); }

// packed fp32 pairs (sm_100): two IEEE adds / fmas in one issue slot, bit-identical to the scalar instructions
__device__ __forceinline__ void add_f32x2(float& a0, float& a1, float b0, float b1) {
    asm("{\n\t.reg .b64 x, y;\n\tmov.b64 x, {%0, %1};\n\tmov.b64 y, {%2, %3};\n\tadd.rn.f32x2 x, x, y;\n\tmov.b64 {%0, %1}, x;\n\t}"
        : "+f"(a0), "+f"(a1) : "f"(b0), "f"(b1));
}
__device__ __forceinline__ void fma_f32x2(float& a0, float& a1, float m, float c) {      // a = m * a + c on both
    asm("{\n\t.reg .b64 x, y, z;\n\tmov.b64 x, {%0, %1};\n\tmov.b64 y, {%2, %2};\n\tmov.b64 z, {%3, %3};\n\tfma.rn.f32x2 x, y, x, z;\n\tmov.b64 {%0, %1}, x;\n\t}"
        : "+f"(a0), "+f"(a1) : "f"(m), "f"(c));
}

__device__ __forceinline__ void mul_f32x2(float& a0, float& a1, float b0, float b1) {
    asm("{\n\t.reg .b64 x, y;\n\tmov.b64 x, {%0, %1};\n\tmov.b64 y, {%2, %3};\n\tmul.rn.f32x2 x, x, y;\n\tmov.b64 {%0, %1}, x;\n\t}"
        : "+f"(a0), "+f"(a1) : "f"(b0), "f"(b1));
}
// c' = f c + i g for two cells, the rounding order of cell_update()
__device__ __forceinline__ void cell_update_x2(float& c0, float& c1, float f0, float f1, float i0, float i1, float g0, float g1) {
    asm("{\n\t.reg .b64 c, f, i, g;\n\tmov.b64 c, {%0, %1};\n\tmov.b64 f, {%2, %3};\n\tmov.b64 i, {%4, %5};\n\tmov.b64 g, {%6, %7};\n\t"
        "mul.rn.f32x2 i, i, g;\n\tfma.rn.f32x2 c, f, c, i;\n\tmov.b64 {%0, %1}, c;\n\t}"
        : "+f"(c0), "+f"(c1) : "f"(f0), "f"(f1), "f"(i0), "f"(i1), "f"(g0), "f"(g1));
}

#define AVC_TRACE2(g, pt) do { if (TRACE && tr_cta && (unsigned)(t - TRACE_T0) < (unsigned)TRACE_STEPS) \
    a.trace[((t - TRACE_T0) * 2 + (g)) * TR_PTS + (pt)] = clock64(); } while (0)

// TRACE: the instantiation AVC_LSTM_TRACE=1 launches (clock64 stamps of CTA (0,0)); the production one carries none of it -
// every instruction on a warp's path through a step costs ~9 cycles of the step (profiles/r2_ll_lstm_notes.md).
template <int NKB, bool GXH, int NSUB, bool TRACE>
__global__ void __launch_bounds__(threads(NSUB, NKB), 1)
ll_lstm_kernel(const __grid_constant__ CUtensorMap tmW, const LlLstmArgs a) {
    constexpr int GROUP_WARPS = group_warps(NSUB, NKB);
    constexpr int CPT2 = SUB_N * 4 / GROUP_WARPS;             // accumulator columns per epilogue thread: 8 or 16
    constexpr int ISSUE_WARP = NSUB * GROUP_WARPS;            // and ISSUE_WARP + 1: one K half each
    constexpr uint32_t D_COLS2 = 2 * NSUB * SUB_N;            // accumulator of (sub-tile g, K half ih) at columns (2 g + ih) * 32
    constexpr int MAX_TMEM_CHUNKS2 = max_tmem_chunks(NSUB);
    constexpr int NKB_T = NKB < MAX_TMEM_CHUNKS2 ? NKB : MAX_TMEM_CHUNKS2;
    constexpr int NKB_S = NKB - NKB_T;
    // WHICH chunks live in shared memory: the first HS of each K half (chunks 0.., NKB/2..), i.e. the start of each issuer's first
    // part.  MMAs with A in shared memory take 40 cycles against 17: issued first, they run while the later parts are still
    // arriving; kept at the end of K (the one-tile kernel's choice) they would be the 650-cycle tail behind the last arrival.
    constexpr int HS = NKB_S / 2, HT = NKB / 2 - HS;          // per K half: chunks in shared memory / in tensor memory
    static_assert(NKB_S % 2 == 0 && HS <= NKB / 4, "smem-resident chunks must fit the first quarter of each K half");
    constexpr int PSTRIDE = SUB_N * PLANE_ELEMS;          // elements between planes of a 32-row tile of Yx
    static_assert(NKB == 16 || NKB == 8 || NKB == 4, "H must be 1024, 512 or 256");

    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* sW = smem;                                        // NKB_S x 16 KiB
    uint8_t* sH = sW + NKB_S * W_BLK_BYTES;                    // [NSUB sub-tiles][NKB][2 KiB]
    constexpr int GXB = gx2_bytes(GXH);
    uint8_t* sGx = sH + NSUB * NKB * HBLK2;                    // [NSUB sub-tiles][2 buffers][GXB]: Gx_t is staged ONE STEP AHEAD
    uint8_t* sGT = sGx + 2 * NSUB * GXB;                       // [NSUB][16 KiB]
    uint64_t* bars = reinterpret_cast<uint64_t*>(sGT + NSUB * GT_BYTES);
    uint64_t* wfull = bars;                                    // [1]
    uint64_t* acc_full = bars + 1;                             // [2] the MMAs (both K halves) of a sub-tile's step are complete
    uint64_t* gfull = bars + 4;                                // [2][PARTS] part p of sub-tile g's K chunks is in shared memory (both CTAs)
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 20);
    uint64_t* wst_full = bars + 22;                            // [WST] prologue: a 16 KiB chunk of W_hh has landed in staging buffer i
    uint64_t* wst_empty = bars + 26;                           // [WST] ... and has been moved on to tensor memory
    // staging buffers of that prologue: as many 16 KiB chunks in flight as the not yet used operand / Gx / transpose areas hold, up
    // to four (two in flight filled tensor memory at 12 B/clk: 7 of the launch's 19 fixed us)
    constexpr int WST_ROOM = NSUB * (NKB * HBLK2 + 2 * gx2_bytes(GXH) + GT_BYTES) / W_BLK_BYTES;
    constexpr int WST = WST_ROOM < 4 ? WST_ROOM : 4;
    static_assert(WST >= 2, "the W_hh staging ring needs two buffers");
    // K is handed to the issuers in PARTS pieces.  One sub-tile: eight, each the chunks of two loader warps of either CTA - the
    // MMAs left to issue when the last piece of h_{t-1} lands are an eighth of the step's (same-box A/B: config 3 9.75 -> 9.57 ms,
    // B = 1 p50 1.18 -> 1.16 ms).  Two sub-tiles (and H = 256): quarters - there the issuers share their schedulers with the other
    // sub-tile's epilogue, and twice the barrier waits cost more issue slots than the shorter tail returns (lstm2 0.875 -> 0.90 ms).
    constexpr int PARTS = (NSUB == 1 && NKB >= 8) ? 8 : 4;
    constexpr int WPP = GROUP_WARPS / PARTS;                           // loader warps per part in each CTA
    constexpr int CPP = NKB / PARTS;                                   // chunks per part: 4, 2 or 1

    const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0), lane = threadIdx.x & 31;
    const int ns = blockIdx.x;            // unit slice (32 units)
    const bool tr_cta = TRACE && a.trace != nullptr && blockIdx.x == 0 && blockIdx.y == 0;      // AVC_LSTM_TRACE=1: CTA (0,0) stamps its steps
    const int bt = blockIdx.y;            // group of NSUB 32-utterance sub-tiles
    const uint32_t crank = ptx::cluster_ctarank();
    const size_t tile_elems = (size_t)a.H * SUB_N;                      // one (t, 32-row tile) block of Yx
    const size_t step_elems = tile_elems * a.tiles;

#define AVC_TRACE_PRO(pt) do { if (TRACE && a.trace && blockIdx.x == 0 && blockIdx.y == 0 && threadIdx.x == 0) a.trace[2 * TRACE_STEPS * TR_PTS + (pt)] = clock64(); } while (0)
    AVC_TRACE_PRO(0);
    if (warp == 0) {
        if (lane == 0) {
            ptx::prefetch_tmap(&tmW);
            ptx::mbar_init(wfull, 1);
            for (int i = 0; i < 2 * PARTS; ++i) ptx::mbar_init(&gfull[i], 2 * WPP);          // the part's loader warps of both CTAs
            ptx::mbar_init(&acc_full[0], 2);                                                 // one commit per issuing warp
            ptx::mbar_init(&acc_full[1], 2);
            for (int i = 0; i < WST; ++i) { ptx::mbar_init(&wst_full[i], 1); ptx::mbar_init(&wst_empty[i], 4); }
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
    AVC_TRACE_PRO(1);

    // ---- resident W: tensor-memory slot s holds chunk (s / HT) * NKB/2 + HS + s % HT, shared-memory slot j chunk (j / HS) * NKB/2 + j % HS ----
    // Tensor-memory chunks travel global -> (TMA, 16 KiB swizzled boxes, two staging buffers in the not yet used operand-tile /
    // Gx area) -> registers (thread = gate row: eight conflict-free 16-byte reads) -> tcgen05.st.  One row per thread straight
    // from global memory (16 bytes per lane at a 2 KiB stride, a dependent round per chunk) took 7 of the launch's 19 fixed us.
    {
        uint8_t* wstage = sH;                                  // WST x 16 KiB (sH .. end of the Gx / transpose areas)
        if (warp < 4) {
            const int R = warp * 32 + lane;
            const uint32_t wl = tmem_base + W_COL + ((uint32_t)(warp * 32) << 16);
            for (int sl = 0; sl < NKB_T; ++sl) {
                const int st = sl % WST;
                ptx::mbar_wait(&wst_full[st], (uint32_t)(sl / WST) & 1);
                const uint32_t row = ptx::smem_u32(wstage) + (uint32_t)(st * W_BLK_BYTES + R * 128);
                uint4 v[8];
#pragma unroll
                for (int i = 0; i < 8; ++i) v[i] = ld_shared_v4(row + (uint32_t)((i ^ (R & 7)) << 4));
                ptx::tmem_st32(wl + (uint32_t)(sl * 32), reinterpret_cast<const uint32_t(&)[32]>(v));
                __syncwarp();
                if (lane == 0) ptx::mbar_arrive(&wst_empty[st]);       // the reads above have returned (the store consumed them)
            }
            ptx::tmem_st_wait();
        } else if (warp == 4 && lane == 0) {
            if (NKB_S > 0) {                                   // the shared-memory-resident chunks first: they need no staging buffer
                ptx::mbar_expect_tx(wfull, (uint32_t)(NKB_S * W_BLK_BYTES));
                for (int j = 0; j < NKB_S; ++j)
                    ptx::tma_load_2d(sW + j * W_BLK_BYTES, &tmW, wfull, ((j / (HS > 0 ? HS : 1)) * (NKB / 2) + j % (HS > 0 ? HS : 1)) * KBLK, ns * MROWS);
            }
            for (int sl = 0; sl < NKB_T; ++sl) {
                const int st = sl % WST, kc = (sl / HT) * (NKB / 2) + HS + sl % HT;
                if (sl >= WST) ptx::mbar_wait(&wst_empty[st], (uint32_t)(sl / WST - 1) & 1);
                ptx::mbar_expect_tx(&wst_full[st], (uint32_t)W_BLK_BYTES);
                ptx::tma_load_2d(wstage + st * W_BLK_BYTES, &tmW, &wst_full[st], kc * KBLK, ns * MROWS);
            }
            if (NKB_S > 0) ptx::mbar_wait(wfull, 0);
        }
    }
    AVC_TRACE_PRO(2);
    ptx::tc_fence_before();
    __syncthreads();
    ptx::cluster_sync();
    ptx::tc_fence_after();
    AVC_TRACE_PRO(3);

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
                        // a sub-tile with <= 8 utterances: N = 16 MMAs (each CTA supplies rows 0..7 of its tile, the leader's are
                        // utterances 0..7 = accumulator columns 0..7; the other columns belong to no utterance) - half the pipe time
                        const uint32_t idesc = a.B - (bt * NSUB + g) * SUB_N <= 8 ? a.idesc16 : a.idesc;
                        constexpr uint32_t d_tmem = tmem_base + (uint32_t)((2 * g + ih) * SUB_N);
                        // Issue loop shape: ROLLED over the K chunks with running descriptor words.  ptxas then keeps the whole
                        // sequence on the uniform datapath (UIADD3 / UMOV between the UTCHMMAs); unrolled, it computes every
                        // descriptor in vector registers and pays ~2 R2UR per MMA - 27 cycles per MMA from one warp, against
                        // the 17 the pipe needs at N = 32 (tools/micro/mma_burst.cu, profiles/r2_ll_lstm_notes.md).
                        auto issue_part = [&](auto P) {
                            constexpr int p = decltype(P)::value;
                            constexpr int c0 = p * CPP, c1 = c0 + CPP;
                            constexpr int ss_lim = ih * (NKB / 2) + HS;                   // end of this K half's smem-resident chunks
                            constexpr int ss_end = ss_lim < c0 ? c0 : (ss_lim > c1 ? c1 : ss_lim);   // [c0, ss_end) A in smem, [ss_end, c1) A in TMEM
                            ptx::mbar_wait(&gfull[g * PARTS + p], par);
                            ptx::tc_fence_after();
                            if (p == 0 && lane == 0) AVC_TRACE2(g, 8);
                            if (p % (PARTS / 4) == 0 && lane == 0) AVC_TRACE2(g, 10 + p / (PARTS / 4));
                            const uint64_t bd0 = ptx::umma_desc_sw128(sH_u + (uint32_t)((g * NKB + c0) * HBLK2));
                            const uint32_t b_hi = (uint32_t)(bd0 >> 32);
                            uint32_t b_lo = (uint32_t)bd0;
                            if constexpr (c0 < ss_end) {
                                const uint64_t ad0 = ptx::umma_desc_sw128(sW_u + (uint32_t)((ih * HS + c0 - ih * (NKB / 2)) * W_BLK_BYTES));
                                const uint32_t a_hi = (uint32_t)(ad0 >> 32);
                                uint32_t a_lo = (uint32_t)ad0;
#pragma unroll 1
                                for (int c = c0; c < ss_end; ++c) {
#pragma unroll
                                    for (int k = 0; k < KBLK / 16; ++k) {
                                        ptx::umma2_f16_warp(d_tmem, ((uint64_t)a_hi << 32) | a_lo, ((uint64_t)b_hi << 32) | b_lo, idesc,
                                                            (c != ih * (NKB / 2) || k != 0) ? 1u : 0u);
                                        a_lo += 2; b_lo += 2;
                                    }
                                    a_lo += (uint32_t)(W_BLK_BYTES >> 4) - 2 * (KBLK / 16);
                                    b_lo += (uint32_t)(HBLK2 >> 4) - 2 * (KBLK / 16);
                                }
                            }
                            if constexpr (ss_end < c1) {
                                // tensor-memory slot of chunk ss_end: ih * HT + (ss_end - ih * NKB/2 - HS)
                                uint32_t a_t = tmem_base + W_COL + (uint32_t)((ih * HT + ss_end - ih * (NKB / 2) - HS) * 32);
#pragma unroll 1
                                for (int c = ss_end; c < c1; ++c) {
#pragma unroll
                                    for (int k = 0; k < KBLK / 16; ++k) {
                                        ptx::umma2_f16_ts_warp(d_tmem, a_t, ((uint64_t)b_hi << 32) | b_lo, idesc, (c != ih * (NKB / 2) || k != 0) ? 1u : 0u);
                                        a_t += 8; b_lo += 2;
                                    }
                                    b_lo += (uint32_t)(HBLK2 >> 4) - 2 * (KBLK / 16);
                                }
                            }
                        };
                        constexpr int P0 = ih * (PARTS / 2);
                        issue_part(std::integral_constant<int, P0>{});
                        issue_part(std::integral_constant<int, P0 + 1>{});
                        if constexpr (PARTS == 8) {
                            issue_part(std::integral_constant<int, P0 + 2>{});
                            issue_part(std::integral_constant<int, P0 + 3>{});
                        }
                        ptx::umma2_commit_warp(&acc_full[g]);       // -> acc_full[g] of BOTH CTAs
                        if (ih == 1 && lane == 0) AVC_TRACE2(g, 9);
                    };
                    issue_sub(std::integral_constant<int, 0>{});
                    if constexpr (NSUB == 2) issue_sub(std::integral_constant<int, 1>{});
                }
            };
            if (warp == ISSUE_WARP) run(std::integral_constant<int, 0>{}); else run(std::integral_constant<int, 1>{});
        }
    } else {
        // ===================== sub-tile group g: loaders + epilogue =====================
        const int g = warp / GROUP_WARPS, wg = warp % GROUP_WARPS;
        const int tg = (int)(threadIdx.x & ((GROUP_WARPS * 32) - 1));
        const int b0 = (bt * NSUB + g) * SUB_N;
        __nv_bfloat16* yx_tile = a.Yx + (size_t)(a.tile0 + bt * NSUB + g) * tile_elems;    // + t * step_elems
        uint64_t* my_acc = &acc_full[g];
        // ---- loader geometry: warp wg owns chunks 2wg, 2wg+1; piece i: chunk 2wg + (i>>2), plane 2(i&3) + (lane>>4), row lane&15 ----
        const int lrow = lane & 15, lpl = lane >> 4;
        const uint32_t sH_g = ptx::smem_u32(sH) + (uint32_t)(g * NKB * HBLK2);
        constexpr int CPW = NKB / GROUP_WARPS > 0 ? NKB / GROUP_WARPS : 1;      // chunks per warp: 2 (H=1024), 1 (H=512)
        constexpr int WPC2 = NKB >= GROUP_WARPS ? 1 : GROUP_WARPS / NKB;          // warps per chunk: 1, or 2 (H=256)
        constexpr int NP = SUB_ROWS * 8 * NKB / (GROUP_WARPS * 32);                  // 16-byte pieces per lane and step: 8 / 4 / 2
        auto piece_chunk = [&](int i) { return WPC2 == 1 ? wg * CPW + (i >> 2) : wg / WPC2; };
        auto piece_plane = [&](int i) { return WPC2 == 1 ? 2 * (i & 3) + lpl : (wg % WPC2) * (8 / WPC2) + 2 * i + lpl; };
        const __nv_bfloat16* rows_base = yx_tile + (size_t)(crank * SUB_ROWS + lrow) * PLANE_ELEMS;
        // The lane's pieces sit at compile-time distances from its first one; everything that does not depend on t is computed
        // here, once, and made opaque to the compiler (it would rather rebuild 64-bit addresses every step than hold them).
        const char* rd_base = reinterpret_cast<const char*>(rows_base + (size_t)(piece_chunk(0) * 8 + piece_plane(0)) * PSTRIDE);
        asm volatile("" : "+l"(rd_base));
        auto piece_delta = [](int i) { return (WPC2 == 1 ? (i >> 2) * 8 + 2 * (i & 3) : 2 * i) * PSTRIDE * 2; };      // bytes from piece 0
        const uint32_t slot_bytes = (uint32_t)(step_elems * 2);      // one ring slot (<= a few MiB)
        const int gp = wg / WPP;                                // the part this warp's chunks belong to
        uint32_t gfull_remote = ptx::mapa_u32(ptx::smem_u32(&gfull[g * PARTS + gp]), 0u);    // the LEADER's barrier of this warp's part
        asm volatile("" : "+r"(gfull_remote));

        // ---- epilogue geometry ----
        const int q = wg & 3, half = wg >> 2;                 // TMEM lane quarter (= warp % 4); block of CPT2 columns (utterances)
        const int R = q * 32 + lane;                          // gate row of this CTA = TMEM lane
        const int gj = lane & 3;                              // gate of this lane: 0=i 1=f 2=g 3=o
        uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(2 * g * SUB_N + half * CPT2);   // K half 1: + SUB_N
        asm volatile("" : "+r"(taddr));                       // held in a register (else: rebuilt after every accumulator wait)
        const float act_s = gj == 2 ? 2.0f : 1.0f, act_o = gj == 2 ? -1.0f : 0.0f;
        const uint32_t gt_smem = ptx::smem_u32(sGT) + (uint32_t)(g * GT_BYTES + wg * (CPT2 * 32 * 4));   // this warp's transpose tile
        float c[CPT2 / 4];
#pragma unroll
        for (int k = 0; k < CPT2 / 4; ++k) c[k] = 0.0f;
        const uint32_t gx_smem0 = ptx::smem_u32(sGx) + (uint32_t)(g * 2 * GXB);
        // Gx_t staging, one step ahead: the block of step t is copied into buffer t & 1 while step t-1 runs (the copies come from
        // HBM: issued in their own step they would not be back before the MMAs of a 32-wide sub-tile are).  WARP-PRIVATE: a warp
        // copies exactly the [16 utterances of its half] x [32 gate rows of its lane quarter] block its own lanes read, so the
        // hand-over needs a __syncwarp, not a barrier of the group.  One commit group per step (an empty one past the end), so
        // "all but the newest group" is always "this step's block".  The row index t / gx_div is kept incrementally.
        // Source pointers and destination offsets of the lane's pieces are fixed for the launch: only the row offset moves.
        constexpr int GX_ES = GXH ? 2 : 4, GX_SEGS = 32 * GX_ES / 16;     // 16-byte pieces per utterance of the warp's block: 4 / 8
        constexpr int GX_NP = CPT2 * GX_SEGS / 32;                       // pieces per lane and step: 2 / 4
        const char* gx_src[GX_NP];
        uint32_t gx_dst[GX_NP];
#pragma unroll
        for (int i = 0; i < GX_NP; ++i) {
            const int u = half * CPT2 + lane / GX_SEGS + (32 / GX_SEGS) * i, seg = lane % GX_SEGS;
            const int b = b0 + u < a.B ? b0 + u : a.B - 1;
            gx_src[i] = reinterpret_cast<const char*>(a.Gx) + ((size_t)b * a.gx_sb + (size_t)ns * MROWS + (size_t)(q * 32)) * GX_ES + seg * 16;
            gx_dst[i] = gx_smem0 + (uint32_t)((u * MROWS + q * 32) * GX_ES + seg * 16);
        }
        const size_t gx_row_bytes = (size_t)a.gx_st * GX_ES;
        size_t gx_off = 0;                        // byte offset of the row of the NEXT block to stage
        int gx_left = a.gx_div;                   // steps left on that row
        auto stage_gx = [&](int t) {
            if (t < a.T) {
                const uint32_t boff = (uint32_t)((t & 1) * GXB);
#pragma unroll
                for (int i = 0; i < GX_NP; ++i) cp_async16(gx_dst[i] + boff, gx_src[i] + gx_off);
                if (--gx_left == 0) { gx_off += gx_row_bytes; gx_left = a.gx_div; }
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
        // h_t store geometry (see the store block below), fixed for the launch
        const bool wr_odd = (lane >> 2) & 1;
        constexpr int HW = CPT2 / 8;                                                // 32-bit words (utterances) a lane stores per step: 2 or 1
        const int wr_col = half * CPT2 + gj * (CPT2 / 4) + (wr_odd ? HW : 0);       // first of the lane's HW utterances (within the sub-tile)
        const int wr_upair = (lane >> 3) * 2;                                       // even unit of the lane's pair, within this quarter's 8
        char* yx_wr = reinterpret_cast<char*>(yx_tile + ((size_t)(ns * 4 + q) * SUB_N + wr_col) * PLANE_ELEMS + wr_upair);
        char* y_wr = reinterpret_cast<char*>(a.Y + (size_t)ns * SLICE_UNITS + q * 8 + wr_upair + (size_t)(b0 + wr_col < a.B ? b0 + wr_col : 0) * a.y_sb);
        asm volatile("" : "+l"(yx_wr), "+l"(y_wr));
        const bool y_ok0 = b0 + wr_col < a.B, y_ok1 = HW == 2 && b0 + wr_col + 1 < a.B;
        const size_t y_row_bytes = (size_t)a.y_sb * 2, y_step_bytes = (size_t)a.y_st * 2;
        stage_gx(0);
        stage_gx(1);

        for (int t = 0; t < a.T; ++t) {
            gx_smem = gx_smem0 + (uint32_t)((t & 1) * GXB);
            if (t >= 1) {
                // ---- h_{t-1} of this sub-tile: L2 -> registers (poll) -> swizzled operand tile ----
                const char* rd = rd_base + (((t - 1) & 1) ? slot_bytes : 0u);                  // ring slot of h_{t-1} ...
                const uint32_t etag = (((uint32_t)(t - 1) >> 1) & 1u) * EPOCH_BITS;           // ... and its epoch
                if (tg == 0) AVC_TRACE2(g, 0);
                uint4 v[NP];
#pragma unroll
                for (int i = 0; i < NP; ++i) v[i] = ld_strong(rd + piece_delta(i));
                for (uint32_t spin = 0;; ++spin) {          // pieces not written yet: fetch all of them again, together
                    bool missing = false;
#pragma unroll
                    for (int i = 0; i < NP; ++i)
                        if (unwritten(v[i], etag)) { v[i] = ld_strong(rd + piece_delta(i)); missing = true; }
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
                    v[i].x &= ~EPOCH_BITS; v[i].y &= ~EPOCH_BITS; v[i].z &= ~EPOCH_BITS; v[i].w &= ~EPOCH_BITS;      // strip the epoch
                    st_shared_v4(sH_g + (uint32_t)(piece_chunk(i) * HBLK2 + lrow * 128 + ((pl ^ (lrow & 7)) << 4)), v[i]);
                }
                ptx::fence_proxy_async_smem();
                __syncwarp();
                // both CTAs arrive on the leader's barrier.  Default semantics (release at CTA scope): a .release.cluster arrive compiles
                // to MEMBAR.ALL.GPU - an L2 round trip per arrive - and the data it guards is shared memory behind fence.proxy.async
                if (lane == 0) asm volatile("mbarrier.arrive.shared::cluster.b64 _, [%0];" ::"r"(gfull_remote) : "memory");
                if (tg == 0) AVC_TRACE2(g, 2);
                if (lane == 0 && wg < 8) AVC_TRACE2(g, 16 + wg);
                stage_gx(t + 1);
            }
            // ---- epilogue: gates, cell state, h_t ----
            if (tg == 0) AVC_TRACE2(g, 24);
            asm volatile("cp.async.wait_group 1;" ::: "memory");     // own pieces of Gx_t (the newest group is Gx_{t+1}) ...
            __syncwarp();                                            // ... and the other lanes' (the block is this warp's alone)
            if (tg == 0) AVC_TRACE2(g, 26);
            // Gx_t of this lane's row, fetched and converted BEFORE the accumulator wait: 32 instructions off the critical path
            float gxv[CPT2];
#pragma unroll
            for (int k = 0; k < CPT2; ++k) gxv[k] = gx_at(k);
            float act[CPT2];
            if (t >= 1) {
                ptx::mbar_wait(my_acc, (uint32_t)(t - 1) & 1);
                if (tg == 0) AVC_TRACE2(g, 3);
                ptx::tc_fence_after();
                uint32_t acc0[CPT2], acc1[CPT2];
                if constexpr (CPT2 == 16) {
                    ptx::tmem_ld16(taddr, acc0);
                    ptx::tmem_ld16(taddr + (uint32_t)SUB_N, acc1);
                } else {
                    ptx::tmem_ld8(taddr, acc0);
                    ptx::tmem_ld8(taddr + (uint32_t)SUB_N, acc1);
                }
                ptx::tmem_ld_wait();
                ptx::tc_fence_before();
                if (tg == 0) AVC_TRACE2(g, 5);
#pragma unroll
                for (int k = 0; k < CPT2; k += 2) {        // (acc0 + acc1) + Gx, in this order
                    act[k] = __uint_as_float(acc0[k]); act[k + 1] = __uint_as_float(acc0[k + 1]);
                    add_f32x2(act[k], act[k + 1], __uint_as_float(acc1[k]), __uint_as_float(acc1[k + 1]));
                    add_f32x2(act[k], act[k + 1], gxv[k], gxv[k + 1]);
                }
            } else {
#pragma unroll
                for (int k = 0; k < CPT2; ++k) act[k] = 0.0f + gxv[k];
            }
            // gate activation of this lane's row (x' already is the tanh argument, see pack_lstm), written to the warp's transpose
            // tile [column][gate row]: the four gates of a unit are four consecutive floats there
            if (a.approx) {
                const float ps = gj == 2 ? 1.0f : 0.5f, po = gj == 2 ? 0.0f : 0.5f;
#pragma unroll
                for (int k = 0; k < CPT2; k += 2) {
                    asm("tanh.approx.f32 %0, %0;" : "+f"(act[k]));
                    asm("tanh.approx.f32 %0, %0;" : "+f"(act[k + 1]));
                    fma_f32x2(act[k], act[k + 1], ps, po);
                }
            } else {
#pragma unroll
                for (int k = 0; k < CPT2; ++k) act[k] = __fdividef(act_s, 1.0f + __expf(-2.0f * act[k])) + act_o;
            }
            // Transpose tile of the warp, CPT2 / 2 rows (column PAIRS) x 256 bytes: row p holds, per unit u of the lane quarter, the floats
            // [i_k i_k+1 f_k f_k+1 | g_k g_k+1 o_k o_k+1] of columns k = 2p, 2p+1 - a lane (gate row 4u + j) stores its pair as one
            // 8-byte word, a reader (unit, column group) gets every operand of the two cells as an adjacent register pair for the
            // packed fp32x2 instructions.  The 16-byte granules of a row are XOR-swizzled by the row's column group: a quarter-warp
            // reads the rows of four groups at a 512-byte stride - unswizzled a 4-way bank conflict (12.5 M of the launch's 21.5 M
            // shared-load wavefronts were such conflicts).
            static_assert(CPT2 == 16 || CPT2 == 8, "tile geometry below: 4 or 2 columns per lane in the cell stage");
            auto gt_swz = [](int grp) { return (uint32_t)((grp & 1) | ((grp & 2) << 1)); };      // 0, 1, 4, 5
#pragma unroll
            for (int kp = 0; kp < CPT2 / 2; ++kp)
                asm volatile("st.shared.v2.f32 [%0], {%1, %2};" ::"r"(gt_smem + (uint32_t)(kp * 256) + ((((uint32_t)lane >> 1) ^ gt_swz(kp / (CPT2 / 8))) << 4) + (uint32_t)((lane & 1) << 3)),
                             "f"(act[2 * kp]), "f"(act[2 * kp + 1]) : "memory");
            __syncwarp();
            if (tg == 0) AVC_TRACE2(g, 6);
            // lane j: unit (j >> 2) of this quarter, columns 4 (j & 3) .. + 3 of this half
            constexpr int H2 = CPT2 / 4;
            uint32_t hp[H2 / 2];
            const uint32_t gt_rd = gt_smem + (uint32_t)(gj * (H2 / 2) * 256) + (((uint32_t)(lane >> 2) * 2u) ^ gt_swz(gj)) * 16u;      // (i, f) granule; (g, o): ^ 16
#pragma unroll
            for (int k = 0; k < H2; k += 2) {
                const uint4 gif = ld_shared_v4(gt_rd + (uint32_t)((k >> 1) * 256));               // i_k i_k+1 f_k f_k+1
                const uint4 ggo = ld_shared_v4((gt_rd ^ 16u) + (uint32_t)((k >> 1) * 256));        // g_k g_k+1 o_k o_k+1
                cell_update_x2(c[k], c[k + 1], __uint_as_float(gif.z), __uint_as_float(gif.w), __uint_as_float(gif.x), __uint_as_float(gif.y),
                               __uint_as_float(ggo.x), __uint_as_float(ggo.y));
                float hv[2];
#pragma unroll
                for (int w = 0; w < 2; ++w) {
                    if (a.approx) asm("tanh.approx.f32 %0, %1;" : "=f"(hv[w]) : "f"(c[k + w]));
                    else hv[w] = 1.0f - __fdividef(2.0f, 1.0f + __expf(2.0f * c[k + w]));
                }
                mul_f32x2(hv[0], hv[1], __uint_as_float(ggo.z), __uint_as_float(ggo.w));          // h = tanh(c') o
                __nv_bfloat162 h2 = __floats2bfloat162_rn(hv[0], hv[1]);
                hp[k >> 1] = *reinterpret_cast<uint32_t*>(&h2);
            }
            __syncwarp();        // the tile is rewritten by the next step's activations
            if (tg == 0) AVC_TRACE2(g, 7);
            // h_t out, straight from registers.  The lane holds one unit x four utterances (hp[0] = columns 0,1; hp[1] = columns 2,3
            // of its block); lanes j and j ^ 4 hold neighbouring units of the same utterances.  One exchange gives every lane a
            // (unit pair) x (two utterances) block, i.e. two 32-bit words: the validity protocol needs whole 32-bit words from one
            // thread (a word with one stale half must not pass the epoch check).  No shared-memory staging, no barrier.
            {
                // HW = 2: w0 / w1 = the lane's first / second utterance.  HW = 1 (two utterances per lane, 16 warps): the even-unit lane
                // stores the pair's word of the first utterance, the odd-unit lane that of the second.
                uint32_t w0, w1 = 0u;
                if constexpr (HW == 2) {
                    const uint32_t mine = wr_odd ? hp[1] : hp[0];
                    const uint32_t theirs = __shfl_xor_sync(0xffffffffu, wr_odd ? hp[0] : hp[1], 4);
                    const uint32_t ev = wr_odd ? theirs : mine, od = wr_odd ? mine : theirs;      // even unit / odd unit, same two utterances
                    w0 = __byte_perm(ev, od, 0x5410); w1 = __byte_perm(ev, od, 0x7632);
                } else {
                    const uint32_t theirs = __shfl_xor_sync(0xffffffffu, hp[0], 4);
                    const uint32_t ev = wr_odd ? theirs : hp[0], od = wr_odd ? hp[0] : theirs;
                    w0 = __byte_perm(ev, od, wr_odd ? 0x7632 : 0x5410);
                }
                if (t + 1 < a.T) {
                    uint32_t x0 = w0, x1 = w1;
                    if ((x0 | x1) & EPOCH_BITS) {      // non-finite h (never for finite inputs): +-1 in the exchange copy, see the header
                        auto fix = [](uint32_t w) {
                            if (w & 0x00004000u) w = (w & 0xffff8000u) | 0x00003f80u;
                            if (w & 0x40000000u) w = (w & 0x8000ffffu) | 0x3f800000u;
                            return w;
                        };
                        x0 = fix(x0); x1 = fix(x1);
                    }
                    const uint32_t ptag = (((uint32_t)t >> 1) & 1u) * EPOCH_BITS;
                    char* yx = yx_wr + ((t & 1) ? slot_bytes : 0u);
                    asm volatile("st.relaxed.gpu.global.u32 [%0], %1;" ::"l"(yx), "r"(x0 | ptag) : "memory");
                    if constexpr (HW == 2) asm volatile("st.relaxed.gpu.global.u32 [%0], %1;" ::"l"(yx + PLANE_ELEMS * 2), "r"(x1 | ptag) : "memory");
                }
                if (y_ok0) *reinterpret_cast<uint32_t*>(y_wr) = w0;
                if (HW == 2 && y_ok1) *reinterpret_cast<uint32_t*>(y_wr + y_row_bytes) = w1;
                y_wr += y_step_bytes;
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

template <int NSUB, bool TRACE>
const void* kernel_for_n(int nkb, bool gxh) {
    switch (nkb) {
        case 16: return gxh ? (const void*)ll_lstm_kernel<16, true, NSUB, TRACE> : (const void*)ll_lstm_kernel<16, false, NSUB, TRACE>;
        case 8: return gxh ? (const void*)ll_lstm_kernel<8, true, NSUB, TRACE> : (const void*)ll_lstm_kernel<8, false, NSUB, TRACE>;
        default: return gxh ? (const void*)ll_lstm_kernel<4, true, NSUB, TRACE> : (const void*)ll_lstm_kernel<4, false, NSUB, TRACE>;
    }
}
const void* kernel_for(int nkb, bool gxh, int nsub, bool trace = false) {
    if (trace) return nsub == 2 ? kernel_for_n<2, true>(nkb, gxh) : kernel_for_n<1, true>(nkb, gxh);
    return nsub == 2 ? kernel_for_n<2, false>(nkb, gxh) : kernel_for_n<1, false>(nkb, gxh);
}
int smem_for(int nkb, bool gxh, int nsub) {
    const int nkb_s = nkb > max_tmem_chunks(nsub) ? nkb - max_tmem_chunks(nsub) : 0;
    return nkb_s * W_BLK_BYTES + nsub * (nkb * HBLK2 + 2 * gx2_bytes(gxh) + GT_BYTES) + 1024 + 512;
}
constexpr int SMEM_LIMIT = 227 * 1024;

}  // namespace

struct LlLstmPlan {
    CUtensorMap tmW;
    std::vector<LlLstmArgs> args;
    std::vector<dim3> grids;
    int nkb;
    int nsub;                 // 32-utterance sub-tiles per CTA pair: 1 or 2
    bool gx_f16;
    int smem_bytes;
    void* yx_base; size_t yx_bytes;       // the exchange ring, pre-filled with 0xFF before every launch
};

const char* ll_lstm_init() {
    if (g_encode_ll) return nullptr;
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult qres;
    cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres);
    if (e != cudaSuccess || qres != cudaDriverEntryPointSuccess || !fn) return "cuTensorMapEncodeTiled not available from the driver";
    for (int i = 0; i < 12; ++i) {
        const int nkb = 4 << ((i >> 1) % 3), nsub = i < 6 ? 1 : 2;
        const int bytes = smem_for(nkb, i & 1, nsub);
        if (bytes > SMEM_LIMIT) continue;          // H = 1024, two sub-tiles, fp32 Gx does not fit: such a plan takes one sub-tile
        for (int tr = 0; tr < 2; ++tr)
            if (cudaFuncSetAttribute(kernel_for(nkb, i & 1, nsub, tr), cudaFuncAttributeMaxDynamicSharedMemorySize, bytes) != cudaSuccess)
                return "cudaFuncSetAttribute(MaxDynamicSharedMemorySize) failed for ll_lstm_kernel";
    }
    g_encode_ll = reinterpret_cast<EncodeTiledFn>(fn);
    return nullptr;
}

size_t ll_lstm_exchange_bytes(int B, int T, int H) { (void)T; return (size_t)2 * ceil_div(B, 2 * SUB_N) * 2 * SUB_N * H * 2; }   // the two-slot ring

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
    plan->gx_f16 = gx_type == DT_F16;
    // One 32-utterance tile per CTA pair while all tiles fit the machine at once (a step is a latency chain: every tile on its own
    // SMs is the shortest); two pipelined sub-tiles per pair beyond that (lstm2 from B = 129 on).  AVC_LSTM_NSUB=1/2 forces it.
    int nsub = ceil_div(B, SUB_N) * G_n <= (num_sms & ~1) ? 1 : 2;
    if (const char* pv = getenv("AVC_LSTM_NSUB")) nsub = atoi(pv) == 2 ? 2 : 1;
    if (smem_for(nkb, plan->gx_f16, nsub) > SMEM_LIMIT) nsub = 1;    // H = 1024 with fp32 Gx (AVC_GX_F32): no room for two sub-tiles' Gx buffers
    plan->nsub = nsub;
    plan->smem_bytes = smem_for(nkb, plan->gx_f16, nsub);
    const int tnp = SUB_N * nsub;                                    // utterances per CTA pair
    plan->yx_base = Yx; plan->yx_bytes = (size_t)2 * ceil_div(B, tnp) * tnp * H * 2;       // two step slots
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
    const int pairs_per_launch = (num_sms & ~1) / G_n;               // batch tiles (CTA pairs along y) of one co-resident launch
    const int rows_per_launch = pairs_per_launch * tnp;
    for (int bs = 0; bs < B; bs += rows_per_launch) {                // B beyond one co-resident grid: several launches
        const int Bc = (B - bs) < rows_per_launch ? (B - bs) : rows_per_launch;
        LlLstmArgs a{};
        a.B = Bc; a.T = T; a.H = H; a.gx_div = gx_div;
        a.y_sb = (long long)T * H; a.y_st = H; a.gx_sb = (long long)Tg * 4 * H; a.gx_st = 4LL * H;
        a.Gx = (const char*)Gx + (size_t)bs * a.gx_sb * dtype_size(gx_type);
        a.Whh = reinterpret_cast<const __nv_bfloat16*>(Whh_bf16);
        a.Y = reinterpret_cast<__nv_bfloat16*>(Y) + (size_t)bs * a.y_sb;
        a.Yx = reinterpret_cast<__nv_bfloat16*>(Yx);
        a.tiles = ceil_div(B, tnp) * nsub; a.tile0 = bs / tnp * nsub;          // Yx is indexed in 32-row tiles
        a.approx = getenv("AVC_LSTM_EXACT") == nullptr;
        a.idesc = ptx::umma_idesc_f16(1, 2 * MROWS, SUB_N);          // the pair's MMA: M = 256 (128 gate rows per CTA), N = 32
        a.idesc16 = getenv("AVC_LSTM_NO_N16") ? a.idesc : ptx::umma_idesc_f16(1, 2 * MROWS, 16);
        plan->args.push_back(a);
        plan->grids.push_back(dim3(G_n, ceil_div(Bc, tnp)));
    }
    return plan;
}

void ll_lstm_plan_destroy(LlLstmPlan* plan) { delete plan; }

cudaError_t launch_ll_lstm(const LlLstmPlan* plan, cudaStream_t s, int* launches) {
    static long long* d_trace = nullptr;
    const bool tracing = getenv("AVC_LSTM_TRACE") != nullptr;
    if (tracing && !d_trace) cudaMalloc(&d_trace, sizeof(long long) * TRACE_WORDS);
    // the ring starts as "epoch 1 everywhere" (0xFF...), which the epoch-0 steps 0 and 1 read as not written yet
    cudaError_t e = cudaMemsetAsync(plan->yx_base, 0xFF, plan->yx_bytes, s);
    if (e != cudaSuccess) return e;
    for (size_t i = 0; i < plan->args.size(); ++i) {
        LlLstmArgs args = plan->args[i];
        args.trace = tracing ? d_trace : nullptr;
        if (tracing) cudaMemsetAsync(d_trace, 0, sizeof(long long) * TRACE_WORDS, s);
        void* kargs[2] = {(void*)&plan->tmW, (void*)&args};
        // CTAs wait on one another's output: cooperative launch (guaranteed co-residency), with the checked non-cooperative
        // retry of launch_persistent() for profilers (AVC_LSTM_NOCOOP forces it)
        e = launch_persistent(kernel_for(plan->nkb, plan->gx_f16, plan->nsub, tracing), plan->grids[i], dim3(threads(plan->nsub, plan->nkb)), (size_t)plan->smem_bytes, 2,
                              kargs, s, getenv("AVC_LSTM_NOCOOP") != nullptr);
        if (e != cudaSuccess) return e;
        if (launches) ++*launches;
        if (tracing) {   // debugging aid: synchronous dump of the per-step timeline of CTA (0,0)
            long long hbuf[TRACE_WORDS];
            cudaStreamSynchronize(s);
            cudaMemcpy(hbuf, d_trace, sizeof hbuf, cudaMemcpyDeviceToHost);
            fprintf(stderr, "ll_lstm trace H=%d B=%d grid=(%d,%d) nsub=%d: per sub-tile, cycles since its loads were issued (warp 0 of the group unless noted): "
                    "landed published | quarters ready q0..q3 (issuers), committed | Gx wait entered, Gx ready | acc_seen tmem_read gates cell stored | "
                    "publish time of warps 0..7 || absolute start\n", args.H, args.B, plan->grids[i].x, plan->grids[i].y, plan->nsub);
            const long long t00 = hbuf[0];
            { const long long* pr = hbuf + 2 * TRACE_STEPS * TR_PTS;
              fprintf(stderr, "  prologue (cycles since kernel entry, thread 0 of CTA (0,0)): barriers+TMEM alloc %lld, W resident (own warp) %lld, cluster synced %lld; first traced step starts at %lld\n",
                      pr[1] - pr[0], pr[2] - pr[0], pr[3] - pr[0], t00 - pr[0]); }
            for (int st = 0; st < TRACE_STEPS; ++st)
                for (int g = 0; g < plan->nsub; ++g) {
                    const long long* r = hbuf + (st * 2 + g) * TR_PTS;
                    if (!r[0]) continue;
                    const long long z = r[0];
                    fprintf(stderr, "  t=%d sub %d: %5lld %5lld | %5lld %5lld %5lld %5lld, %5lld | %5lld %5lld | %5lld %5lld %5lld %5lld %5lld | %5lld %5lld %5lld %5lld %5lld %5lld %5lld %5lld || %7lld\n",
                            TRACE_T0 + st, g, r[1] - z, r[2] - z, r[10] - z, r[11] - z, r[12] - z, r[13] - z, r[9] - z, r[24] - z, r[26] - z, r[3] - z, r[5] - z, r[6] - z, r[7] - z, r[4] - z,
                            r[16] - z, r[17] - z, r[18] - z, r[19] - z, r[20] - z, r[21] - z, r[22] - z, r[23] - z, z - t00);
                }
        }
    }
    return cudaSuccess;
}

}  // namespace avc
