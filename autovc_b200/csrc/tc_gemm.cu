// Tensor-core implicit-GEMM for the conv1d(k=5) / projection / linear layers (sm_100a):
//   TMA (cp.async.bulk.tensor, 128-byte swizzle) -> shared memory ring -> tcgen05.mma (bf16|fp16 x bf16|fp16 -> fp32
//   accumulators in TMEM, double-buffered) -> tcgen05.ld epilogue with fused bias(+folded BatchNorm) + ReLU/tanh,
//   writing bf16/fp16/fp32 channel-last activations and/or the reference's fp32 channel-first (B,80,T) layout
//   (+ the postnet residual).
//
// The convolution's zero padding (reference net_layers.py:39, padding=2) is the TMA out-of-bounds zero fill:
// activations are a 3-D tensor (C, T, B) and the tile for tap k is fetched at time coordinate t0 + k - 2, so rows
// that fall outside [0,T) of the SAME utterance arrive as zeros and no padded copy is ever materialised.
//
// Warp roles (192 threads): warp 0 = TMA producer, warp 1 = TMEM allocator + MMA issuer (one elected lane),
// warps 2..5 = epilogue (TMEM lane quarter = warp_id % 4).  Persistent over output tiles.
#include <cstdio>
#include <cuda.h>

#include "common.cuh"
#include "launch.h"
#include "ptx.cuh"

namespace avc {

namespace {

constexpr int BLOCK_M = 128;
constexpr int BLOCK_K = 64;                       // 64 x 2 B = one 128-byte swizzle row
constexpr int UMMA_K = 16;
constexpr int A_STAGE_BYTES = BLOCK_M * BLOCK_K * 2;
constexpr int SMEM_BUDGET = 200 * 1024;

struct TcGemmArgs {
    int N;
    int taps, kblocks_per_tap, Cp, pad;
    int gT, gB, TB, BB;                  // tiling geometry: rows = (b, t), tile = BB utterances x TB frames
    int tiles_t, tiles_n, num_tiles;
    int T;                               // true frames per utterance (channel-first output addressing)
    const float* bias;
    const float* row_bias;               // [B][5][N] per-row bias (replaces `bias`): encoder conv0's speaker term by edge class of the frame
    int act;
    void* out; int out_type; int ldo;
    float* out_t; const float* res_t;
    uint32_t idesc;
    int full_tiles;                      // tiles [0, full_tiles) are BLOCK_N wide; the rest are the last round's tiles split in two halves
    uint32_t idesc_half;
    int tma_out;                         // row-major output leaves through shared memory + TMA stores (tmO)
    int ob_t, ob_b;                      // frames x utterances of one epilogue warp's 32 tile rows
    double* stat_partial;                // STATS kernels: per (M tile, epilogue warp) column sums / sums of squares [chunk][2][N]
};

constexpr int ZERO_BIAS_N = 8192;
__device__ float g_zero_bias[ZERO_BIAS_N];        // zero-initialised: the bias of layers that have none
constexpr int OUT_STAGE_BYTES = 32 * 128;         // one epilogue warp's 32 rows x 128 bytes, SWIZZLE_128B
constexpr int OUT_STAGING_BYTES = 4 * 2 * OUT_STAGE_BYTES;   // 4 epilogue warps, double-buffered

template <int BLOCK_N>
struct Cfg {
    static constexpr int W_STAGE_BYTES = BLOCK_N * BLOCK_K * 2;
    static constexpr int STAGE_BYTES = A_STAGE_BYTES + W_STAGE_BYTES;
    static constexpr int STAGES = (SMEM_BUDGET / STAGE_BYTES) > 8 ? 8 : (SMEM_BUDGET / STAGE_BYTES);
    static constexpr int ACC_STRIDE = BLOCK_N <= 128 ? 128 : 256;   // TMEM columns per accumulator stage
    static constexpr uint32_t TMEM_COLS = 2 * ACC_STRIDE;
    static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + OUT_STAGING_BYTES + 1024 /*align*/ + 256 /*barriers*/;
    static_assert(W_STAGE_BYTES % 1024 == 0, "stage buffers must stay 1024-byte aligned for SWIZZLE_128B");
};

__device__ __forceinline__ float fast_tanh(float x) {
    // tanh(x) = 1 - 2/(1+e^{2x}); ex2.approx + fast division: abs error ~1e-6 (tanh.approx.f32 is ~5e-4, too coarse
    // for the 1e-2 end-to-end budget through four postnet layers)
    const float e = __expf(2.0f * x);
    return 1.0f - __fdividef(2.0f, 1.0f + e);
}

// Static schedule: tile = blockIdx.x + k * gridDim.x.  When the last round would leave more than half of the CTAs idle, its
// tiles are split into two half-width tiles each (full_tiles is a multiple of the grid), e.g. 512 tiles on 148 CTAs:
// 3 full rounds + 68 tiles -> 3 rounds + 136 half tiles = 3.5 instead of 4 rounds.
template <int BLOCK_N>
__device__ __forceinline__ void decode_tile(const TcGemmArgs& a, int tile, int& mt, int& n0, int& bn) {
    if (tile < a.full_tiles) {
        mt = tile / a.tiles_n; n0 = (tile % a.tiles_n) * BLOCK_N; bn = BLOCK_N;
    } else {
        const int h = tile - a.full_tiles, f = a.full_tiles + (h >> 1);
        mt = f / a.tiles_n; n0 = (f % a.tiles_n) * BLOCK_N + (h & 1) * (BLOCK_N / 2); bn = BLOCK_N / 2;
    }
}

// STATS: the fp32 row-major epilogue also reduces every staged 32-row block over its rows (per-column sum and sum of squares of
// the VALID rows) for the train-mode BatchNorm (batchnorm.cu): the statistics pass then never re-reads the conv output.
template <int BLOCK_N, bool STATS = false>
__global__ void __launch_bounds__(192, 1)
tc_gemm_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmW, const __grid_constant__ CUtensorMap tmO,
               const __grid_constant__ CUtensorMap tmWh, const TcGemmArgs a) {
    using C = Cfg<BLOCK_N>;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* sA = smem;
    uint8_t* sW = smem + C::STAGES * A_STAGE_BYTES;
    uint8_t* sOut = smem + C::STAGES * C::STAGE_BYTES;
    uint64_t* bars = reinterpret_cast<uint64_t*>(sOut + OUT_STAGING_BYTES);
    uint64_t* full = bars;
    uint64_t* empty = bars + C::STAGES;
    uint64_t* tfull = bars + 2 * C::STAGES;
    uint64_t* tempty = tfull + 2;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty + 2);

    const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0), lane = threadIdx.x & 31;   // warp: provably uniform
    const int num_kb = a.taps * a.kblocks_per_tap;

    if (warp == 0 && lane == 0) {
        ptx::prefetch_tmap(&tmA);
        ptx::prefetch_tmap(&tmW);
        if (a.tma_out) ptx::prefetch_tmap(&tmO);
        if (a.full_tiles < a.num_tiles) ptx::prefetch_tmap(&tmWh);
        for (int i = 0; i < C::STAGES; ++i) { ptx::mbar_init(&full[i], 1); ptx::mbar_init(&empty[i], 1); }
        for (int i = 0; i < 2; ++i) { ptx::mbar_init(&tfull[i], 1); ptx::mbar_init(&tempty[i], 128); }
        ptx::fence_barrier_init();
    }
    if (warp == 1) ptx::tmem_alloc<C::TMEM_COLS>(tmem_slot);
    ptx::tc_fence_before();
    __syncthreads();
    ptx::tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        // ===================== TMA producer (whole warp, uniform operands, one elected lane issues) =====================
        int stage = 0; uint32_t phase = 0;
        for (int tile = blockIdx.x; tile < a.num_tiles; tile += gridDim.x) {
            int mt, n0, bn;
            decode_tile<BLOCK_N>(a, tile, mt, n0, bn);
            const int t0 = (mt % a.tiles_t) * a.TB, b0 = (mt / a.tiles_t) * a.BB;
            const bool half = bn != BLOCK_N;
            for (int tap = 0; tap < a.taps; ++tap) {
                for (int kc = 0; kc < a.kblocks_per_tap; ++kc) {
                    ptx::mbar_wait(&empty[stage], phase ^ 1);
                    ptx::mbar_expect_tx_warp(&full[stage], half ? A_STAGE_BYTES + C::W_STAGE_BYTES / 2 : C::STAGE_BYTES);
                    ptx::tma_load_3d_warp(sA + stage * A_STAGE_BYTES, &tmA, &full[stage], kc * BLOCK_K, t0 + tap - a.pad, b0);
                    ptx::tma_load_2d_warp(sW + stage * C::W_STAGE_BYTES, half ? &tmWh : &tmW, &full[stage], tap * a.Cp + kc * BLOCK_K, n0);
                    if (++stage == C::STAGES) { stage = 0; phase ^= 1; }
                }
            }
        }
    } else if (warp == 1) {
        // ===================== MMA issuer (whole warp, uniform operands, one elected lane issues) =====================
        // an `if (lane == 0)` around the loop makes the compiler wrap every UTCHMMA in an ELECT/R2UR/BRA.U.ANY loop
        int stage = 0; uint32_t phase = 0;
        int as = 0; uint32_t aphase = 0;
        for (int tile = blockIdx.x; tile < a.num_tiles; tile += gridDim.x) {
            ptx::mbar_wait(&tempty[as], aphase ^ 1);
            ptx::tc_fence_after();
            const uint32_t idesc = tile < a.full_tiles ? a.idesc : a.idesc_half;
            const uint32_t d_tmem = tmem_base + (uint32_t)(as * C::ACC_STRIDE);
            for (int kb = 0; kb < num_kb; ++kb) {
                ptx::mbar_wait(&full[stage], phase);
                ptx::tc_fence_after();
                const uint64_t adesc = ptx::umma_desc_sw128(ptx::smem_u32(sA + stage * A_STAGE_BYTES));
                const uint64_t bdesc = ptx::umma_desc_sw128(ptx::smem_u32(sW + stage * C::W_STAGE_BYTES));
#pragma unroll
                for (int k = 0; k < BLOCK_K / UMMA_K; ++k) {
                    // advance 32 bytes (16 elements) along K inside the swizzle row: +2 in the (addr>>4) field
                    ptx::umma_f16_warp(d_tmem, adesc + (uint64_t)(k * 2), bdesc + (uint64_t)(k * 2), idesc, (kb | k) != 0 ? 1u : 0u);
                }
                ptx::umma_commit_warp(&empty[stage]);
                if (++stage == C::STAGES) { stage = 0; phase ^= 1; }
            }
            ptx::umma_commit_warp(&tfull[as]);
            if (++as == 2) { as = 0; aphase ^= 1; }
        }
    } else {
        // ===================== epilogue (warps 2..5) =====================
        const int q = warp & 3;                 // TMEM lane quarter this warp may access
        const int r = q * 32 + lane;            // tile row
        int as = 0; uint32_t aphase = 0;
        for (int tile = blockIdx.x; tile < a.num_tiles; tile += gridDim.x) {
            int mt, n0, bn;
            decode_tile<BLOCK_N>(a, tile, mt, n0, bn);
            const int t = (mt % a.tiles_t) * a.TB + (r % a.TB);
            const int b = (mt / a.tiles_t) * a.BB + (r / a.TB);
            const bool valid = (t < a.gT) && (b < a.gB);
            const size_t m = (size_t)b * a.gT + t;
            // bias row of this tile row: the layer's bias, or (row_bias) the utterance's vector for the frame's edge class - frames
            // whose 5-tap window reaches into the zero padding lack the speaker term of the taps that fall outside
            const float* brow = a.bias;
            if (a.row_bias) {
                const int cls = t == 0 ? 1 : (t == 1 ? 2 : (t == a.gT - 2 ? 3 : (t == a.gT - 1 ? 4 : 0)));
                brow = a.row_bias + ((size_t)(b < a.gB ? b : a.gB - 1) * 5 + cls) * a.N;
            }
            // channel-first addressing uses the true utterance length
            size_t ot_base = 0;
            if (a.out_t) { const size_t bb = m / a.T, tt = m - bb * a.T; ot_base = (bb * a.N) * a.T + tt; }

            // bias of a 16-column group, loaded AHEAD of its use (the first before the accumulator wait): consumed straight after
            // its load it was the epilogue's top stall (ncu source counters: 13 % of the kernel's samples on the first FADD).  A layer
            // without bias reads a zero vector (g_zero_bias), so the load is unconditional.
            float4 bnext[4];
            auto load_bias = [&](int n) {
#pragma unroll
                for (int j = 0; j < 4; ++j) bnext[j] = __ldg(reinterpret_cast<const float4*>(brow + n) + j);
            };
            const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(as * C::ACC_STRIDE);
            // Row-major output through TMA: each warp stages 32 rows x 128 bytes (XOR-swizzled 16-byte pieces: conflict-free for the
            // row-per-lane writes) and one lane hands the box to TMA - full-line global writes instead of 32 rows x 16 B per store
            // instruction.  Two staging buffers per warp; rows beyond the tensor are clipped by the TMA unit.  NG = 16-column
            // groups per staging row (4 for 16-bit outputs, 2 for fp32), unrolled: the bias of group g of the NEXT block is
            // requested as soon as group g of this block has used its registers - a whole block (>= 300 cycles) ahead.
            auto epilogue_tma = [&](auto NG_) {
                constexpr int NG = decltype(NG_)::value;
                constexpr int cw = NG * 16;                                  // columns per 128-byte staging row
                float4 bs[NG][4];
#pragma unroll
                for (int g2 = 0; g2 < NG; ++g2)
#pragma unroll
                    for (int j = 0; j < 4; ++j) bs[g2][j] = 16 * g2 < bn ? __ldg(reinterpret_cast<const float4*>(brow + n0 + 16 * g2) + j) : make_float4(0.f, 0.f, 0.f, 0.f);
                ptx::mbar_wait(&tfull[as], aphase);
                ptx::tc_fence_after();
                const int r0 = q * 32;
                const int ct = (mt % a.tiles_t) * a.TB + r0 % a.TB, cb = (mt / a.tiles_t) * a.BB + r0 / a.TB;
                uint8_t* stg = sOut + q * 2 * OUT_STAGE_BYTES;
                const uint32_t my_row = ptx::smem_u32(stg) + (uint32_t)(lane * 128);
#pragma unroll 1
                for (int c0 = 0; c0 < bn; c0 += cw) {
                    const uint32_t buf = (uint32_t)((c0 / cw) & 1) * OUT_STAGE_BYTES;
                    if (lane == 0) ptx::bulk_wait_group_read<1>();           // the store that last read this buffer is done
                    __syncwarp();
#pragma unroll
                    for (int g2 = 0; g2 < NG; ++g2) {
                        const int sub = 16 * g2;
                        uint32_t rr[16];
                        ptx::tmem_ld16(taddr + (uint32_t)(c0 + sub), rr);
                        ptx::tmem_ld_wait();
                        float v[16];
#pragma unroll
                        for (int j = 0; j < 16; j += 4) {
                            v[j + 0] = __uint_as_float(rr[j + 0]); v[j + 1] = __uint_as_float(rr[j + 1]);
                            v[j + 2] = __uint_as_float(rr[j + 2]); v[j + 3] = __uint_as_float(rr[j + 3]);
                            ptx::add_f32x2(v[j + 0], v[j + 1], bs[g2][j >> 2].x, bs[g2][j >> 2].y);
                            ptx::add_f32x2(v[j + 2], v[j + 3], bs[g2][j >> 2].z, bs[g2][j >> 2].w);
                        }
                        if (c0 + cw + sub < bn) {
#pragma unroll
                            for (int j = 0; j < 4; ++j) bs[g2][j] = __ldg(reinterpret_cast<const float4*>(brow + n0 + c0 + cw + sub) + j);
                        }
                        if (a.act == ACT_RELU) {
#pragma unroll
                            for (int j = 0; j < 16; ++j) v[j] = ptx::relu_nan(v[j]);     // like torch.relu: a NaN stays a NaN (fmaxf would return 0)
                        } else if (a.act == ACT_TANH) {
#pragma unroll
                            for (int j = 0; j < 16; ++j) v[j] = fast_tanh(v[j]);
                        }
                        if constexpr (NG == 2) {
#pragma unroll
                            for (int j = 0; j < 4; ++j) {
                                const uint32_t piece = (uint32_t)(sub / 4 + j);
                                asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(my_row + buf + ((piece ^ (uint32_t)(lane & 7)) << 4)),
                                             "f"(v[4 * j]), "f"(v[4 * j + 1]), "f"(v[4 * j + 2]), "f"(v[4 * j + 3]) : "memory");
                            }
                        } else {
                            uint32_t pk[8];
                            if (a.out_type == DT_BF16) {
#pragma unroll
                                for (int j = 0; j < 8; ++j) { __nv_bfloat162 h2 = __floats2bfloat162_rn(v[2 * j], v[2 * j + 1]); pk[j] = *reinterpret_cast<uint32_t*>(&h2); }
                            } else {
#pragma unroll
                                for (int j = 0; j < 8; ++j) { __half2 h2 = __floats2half2_rn(v[2 * j], v[2 * j + 1]); pk[j] = *reinterpret_cast<uint32_t*>(&h2); }
                            }
#pragma unroll
                            for (int j = 0; j < 2; ++j) {
                                const uint32_t piece = (uint32_t)(sub / 8 + j);
                                asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(my_row + buf + ((piece ^ (uint32_t)(lane & 7)) << 4)),
                                             "r"(pk[4 * j]), "r"(pk[4 * j + 1]), "r"(pk[4 * j + 2]), "r"(pk[4 * j + 3]) : "memory");
                            }
                        }
                    }
                    ptx::fence_proxy_async_smem();
                    __syncwarp();
                    if (lane == 0) {
                        ptx::tma_store_3d(&tmO, stg + buf, n0 + c0, ct, cb);
                        ptx::bulk_commit_group();
                    }
                    if constexpr (STATS) {
                        // lane = column of the staged 32 x 32 fp32 block (swizzled 16-byte pieces: conflict-free), rows in order.
                        // SHIFTED sums: the block's first valid row is subtracted before summing, so the fp32 sums carry the
                        // column's spread, not its offset (E[x^2] - mean^2 over raw fp32 sums loses the variance of a channel whose
                        // |mean| >> std).  Chunk record = 4 floats in the two 8-byte slots: (S1, S2), (shift, valid rows);
                        // bn_finalize re-bases every chunk onto one shift in double precision.
                        const uint32_t vmask = __ballot_sync(0xffffffffu, valid);
                        const uint32_t col = ptx::smem_u32(stg) + buf + (uint32_t)((lane & 3) * 4);
                        auto at_row = [&](int rr2) {
                            float x;
                            asm volatile("ld.shared.f32 %0, [%1];" : "=f"(x) : "r"(col + (uint32_t)(rr2 * 128) + ((((uint32_t)lane >> 2) ^ (uint32_t)(rr2 & 7)) << 4)));
                            return x;
                        };
                        const float shift = vmask ? at_row(__ffs(vmask) - 1) : 0.0f;
                        float ssum = 0.0f, ssq = 0.0f;
#pragma unroll
                        for (int rr2 = 0; rr2 < 32; ++rr2) {
                            const float x = at_row(rr2) - shift;
                            if ((vmask >> rr2) & 1u) { ssum += x; ssq = fmaf(x, x, ssq); }
                        }
                        double* pp = a.stat_partial + ((size_t)(mt * 4 + q) * 2) * a.N + (n0 + c0 + lane);
                        *reinterpret_cast<float2*>(pp) = make_float2(ssum, ssq);
                        *reinterpret_cast<float2*>(pp + a.N) = make_float2(shift, (float)__popc(vmask));
                    }
                }
            };
            if (a.tma_out) {
                if (a.out_type == DT_F32) epilogue_tma(std::integral_constant<int, 2>{}); else epilogue_tma(std::integral_constant<int, 4>{});
            } else {
            load_bias(n0);
            ptx::mbar_wait(&tfull[as], aphase);
            ptx::tc_fence_after();
#pragma unroll 1
            for (int c0 = 0; c0 < bn; c0 += 16) {
                uint32_t rr[16];
                ptx::tmem_ld16(taddr + (uint32_t)c0, rr);
                const float4 bcur[4] = {bnext[0], bnext[1], bnext[2], bnext[3]};
                if (c0 + 16 < bn) load_bias(n0 + c0 + 16);
                ptx::tmem_ld_wait();
                float v[16];
                const int n = n0 + c0;
#pragma unroll
                for (int j = 0; j < 16; j += 4) {
                    const float4 bv = bcur[j >> 2];
                    v[j + 0] = __uint_as_float(rr[j + 0]) + bv.x;
                    v[j + 1] = __uint_as_float(rr[j + 1]) + bv.y;
                    v[j + 2] = __uint_as_float(rr[j + 2]) + bv.z;
                    v[j + 3] = __uint_as_float(rr[j + 3]) + bv.w;
                }
                if (a.act == ACT_RELU) {
#pragma unroll
                    for (int j = 0; j < 16; ++j) v[j] = v[j] < 0.0f ? 0.0f : v[j];     // like torch.relu: a NaN stays a NaN (fmaxf would return 0)
                } else if (a.act == ACT_TANH) {
#pragma unroll
                    for (int j = 0; j < 16; ++j) v[j] = fast_tanh(v[j]);
                }
                if (valid) {
                    if (a.out) {
                        if (a.out_type == DT_F32) {
                            float4* o = reinterpret_cast<float4*>(reinterpret_cast<float*>(a.out) + m * a.ldo + n);
#pragma unroll
                            for (int j = 0; j < 4; ++j) o[j] = make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
                        } else {
                            uint32_t pk[8];
                            if (a.out_type == DT_BF16) {
#pragma unroll
                                for (int j = 0; j < 8; ++j) { __nv_bfloat162 h2 = __floats2bfloat162_rn(v[2 * j], v[2 * j + 1]); pk[j] = *reinterpret_cast<uint32_t*>(&h2); }
                            } else {
#pragma unroll
                                for (int j = 0; j < 8; ++j) { __half2 h2 = __floats2half2_rn(v[2 * j], v[2 * j + 1]); pk[j] = *reinterpret_cast<uint32_t*>(&h2); }
                            }
                            uint4* o = reinterpret_cast<uint4*>(reinterpret_cast<uint16_t*>(a.out) + m * a.ldo + n);
                            o[0] = make_uint4(pk[0], pk[1], pk[2], pk[3]);
                            o[1] = make_uint4(pk[4], pk[5], pk[6], pk[7]);
                        }
                    }
                    if (a.out_t) {
                        // residual first, all 16 loads in flight (read-only path: `mel` was written by an earlier kernel); interleaved
                        // with the stores they serialise behind one another - the compiler must assume out_t and res_t alias -
                        // and the epilogue of the last postnet layer took 33 us per tile
                        float rv[16];
#pragma unroll
                        for (int j = 0; j < 16; ++j) rv[j] = a.res_t ? __ldg(a.res_t + ot_base + (size_t)(n + j) * a.T) : 0.0f;
#pragma unroll
                        for (int j = 0; j < 16; ++j) a.out_t[ot_base + (size_t)(n + j) * a.T] = v[j] + rv[j];
                    }
                }
            }
            }
            ptx::tc_fence_before();
            ptx::mbar_arrive(&tempty[as]);
            if (++as == 2) { as = 0; aphase ^= 1; }
        }
    }

    if (a.tma_out && warp >= 2 && lane == 0) ptx::bulk_wait_group<0>();   // staged output has left shared memory
    ptx::tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        ptx::tc_fence_after();
        ptx::tmem_dealloc<C::TMEM_COLS>(tmem_base);
    }
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
EncodeTiledFn g_encode = nullptr;

}  // namespace

struct TcGemmPlan {
    CUtensorMap tmA, tmW, tmO, tmWh;
    TcGemmArgs args;
    int block_n;
    int grid;
    int stat_chunks;                     // > 0: the STATS kernel writes this many (M tile x epilogue warp) chunks of column sums
};

const char* tc_init() {
    if (g_encode) return nullptr;
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult qres;
    cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres);
    if (e != cudaSuccess || qres != cudaDriverEntryPointSuccess || !fn) return "cuTensorMapEncodeTiled not available from the driver";
    g_encode = reinterpret_cast<EncodeTiledFn>(fn);
    cudaFuncSetAttribute(tc_gemm_kernel<256>, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg<256>::SMEM_BYTES);
    cudaFuncSetAttribute(tc_gemm_kernel<128>, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg<128>::SMEM_BYTES);
    cudaFuncSetAttribute(tc_gemm_kernel<256, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg<256>::SMEM_BYTES);
    cudaFuncSetAttribute(tc_gemm_kernel<128, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg<128>::SMEM_BYTES);
    cudaFuncSetAttribute(tc_gemm_kernel<80>, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg<80>::SMEM_BYTES);
    cudaFuncSetAttribute(tc_gemm_kernel<64>, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg<64>::SMEM_BYTES);
    cudaFuncSetAttribute(tc_gemm_kernel<32>, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg<32>::SMEM_BYTES);
    if (cudaGetLastError() != cudaSuccess) return "cudaFuncSetAttribute(MaxDynamicSharedMemorySize) failed";
    return nullptr;
}

TcGemmPlan* tc_gemm_plan_create(const GemmParams& p, int num_sms, const char** err) {
    static thread_local char msg[256];
    auto bad = [&](const char* m) -> TcGemmPlan* { if (err) *err = m; return nullptr; };
    if (!g_encode) return bad("tc_init() not called");
    if (p.a_type != p.w_type || (p.a_type != DT_BF16 && p.a_type != DT_F16)) return bad("operands must both be bf16 or both fp16");
    if ((p.lda * 2) % 16 || (p.ldw * 2) % 16) return bad("row pitches must be multiples of 16 bytes");
    if (p.out && ((p.ldo * dtype_size(p.out_type)) % 16)) return bad("output pitch must be a multiple of 16 bytes");
    if (p.Cp % BLOCK_K) {
        // plain GEMMs may have a ragged K tail: TMA zero-fills both operands beyond Kp
        if (p.taps != 1) return bad("conv Cp must be a multiple of 64");
    }
    if (p.row_bias && (p.taps != KTAPS || p.T < 4)) return bad("row_bias is defined for 5-tap convolutions over utterances of at least 4 frames");
    TcGemmPlan* plan = new TcGemmPlan();
    TcGemmArgs& a = plan->args;
    a.N = p.N; a.taps = p.taps; a.Cp = p.Cp; a.pad = p.taps / 2;
    a.kblocks_per_tap = ceil_div(p.taps == 1 ? p.K : p.Cp, BLOCK_K);
    a.T = p.T;
    a.bias = p.bias; a.row_bias = p.row_bias;
    if (!a.bias && !a.row_bias) {            // the epilogue loads its bias unconditionally: a layer without one adds zeros
        void* z = nullptr;
        if (p.N > ZERO_BIAS_N || cudaGetSymbolAddress(&z, g_zero_bias) != cudaSuccess) {
            delete plan;
            return bad("no zero bias vector for this layer (N too large or cudaGetSymbolAddress failed)");
        }
        a.bias = static_cast<const float*>(z);
    } a.act = p.act; a.out = p.out; a.out_type = p.out_type; a.ldo = p.ldo; a.out_t = p.out_t; a.res_t = p.res_t;

    // ---- M tiling: tile = BB utterances x TB frames, TB*BB = 128 ----
    if (p.taps == 1) { a.gT = p.M; a.gB = 1; a.TB = BLOCK_M; a.BB = 1; }
    else {
        a.gT = p.T; a.gB = p.M / p.T;
        long best = -1; int bestTB = BLOCK_M;
        for (int TB = BLOCK_M; TB >= 8; TB >>= 1) {
            const int BB = BLOCK_M / TB;
            const long cost = (long)ceil_div(a.gT, TB) * TB * ceil_div(a.gB, BB) * BB;
            if (best < 0 || cost < best) { best = cost; bestTB = TB; }
        }
        a.TB = bestTB; a.BB = BLOCK_M / bestTB;
    }
    a.tiles_t = ceil_div(a.gT, a.TB);
    const int tiles_m = a.tiles_t * ceil_div(a.gB, a.BB);

    // ---- N tiling ----
    int bn;
    // widest tile that still gives every SM a tile; small batches (latency mode) take narrow tiles: a CTA's time is what it must
    // take in (A tile + W tile per K block at ~42 B/clk per SM), so more, narrower CTAs shorten the layer although A is re-read
    if (p.N % 256 == 0 && (long)tiles_m * (p.N / 256) >= num_sms) bn = 256;
    else if (p.N % 128 == 0 && ((long)tiles_m * (p.N / 128) >= num_sms / 2 || getenv("AVC_GEMM_NO_NARROW"))) bn = 128;
    else if (p.N % 64 == 0 && (long)tiles_m * (p.N / 64) >= num_sms / 2) bn = 64;
    else if (p.N % 32 == 0 && p.N >= 128) bn = 32;
    else if (p.N % 128 == 0) bn = 128;
    else if (p.N == 80) bn = 80;
    else { delete plan; snprintf(msg, sizeof msg, "N=%d is not a multiple of 128 (or 80)", p.N); return bad(msg); }
    plan->block_n = bn;
    a.tiles_n = p.N / bn;
    a.num_tiles = tiles_m * a.tiles_n;
    plan->grid = a.num_tiles < num_sms ? a.num_tiles : num_sms;
    a.full_tiles = a.num_tiles;
    {   // split the tiles of a mostly idle last round into half-width tiles (see decode_tile)
        const int rem = a.num_tiles % plan->grid;
        if (bn == 256 && a.num_tiles > plan->grid && rem > 0 && 2 * rem <= plan->grid && !getenv("AVC_GEMM_NO_TAIL_SPLIT")) {
            a.full_tiles = a.num_tiles - rem;
            a.num_tiles = a.full_tiles + 2 * rem;
        }
    }
    a.idesc_half = ptx::umma_idesc_f16(p.a_type == DT_BF16 ? 1 : 0, BLOCK_M, bn / 2 >= 16 ? bn / 2 : 16);
    a.idesc = ptx::umma_idesc_f16(p.a_type == DT_BF16 ? 1 : 0, BLOCK_M, bn);

    // ---- tensor maps ----
    const CUtensorMapDataType dt = p.a_type == DT_BF16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT16;
    {
        const cuuint64_t kdim = (cuuint64_t)(p.taps == 1 ? p.K : p.Cin);
        cuuint64_t gdim[3] = {kdim, (cuuint64_t)a.gT, (cuuint64_t)a.gB};
        cuuint64_t gstr[2] = {(cuuint64_t)p.lda * 2, (cuuint64_t)a.gT * p.lda * 2};
        cuuint32_t box[3] = {(cuuint32_t)BLOCK_K, (cuuint32_t)a.TB, (cuuint32_t)a.BB};
        cuuint32_t est[3] = {1, 1, 1};
        CUresult r = g_encode(&plan->tmA, dt, 3, const_cast<void*>(p.A), gdim, gstr, box, est, CU_TENSOR_MAP_INTERLEAVE_NONE,
                              CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (r != CUDA_SUCCESS) { delete plan; snprintf(msg, sizeof msg, "cuTensorMapEncodeTiled(A) failed: %d", (int)r); return bad(msg); }
    }
    {
        cuuint64_t gdim[2] = {(cuuint64_t)p.ldw, (cuuint64_t)p.N};
        cuuint64_t gstr[1] = {(cuuint64_t)p.ldw * 2};
        cuuint32_t box[2] = {(cuuint32_t)BLOCK_K, (cuuint32_t)bn};
        cuuint32_t est[2] = {1, 1};
        CUresult r = g_encode(&plan->tmW, dt, 2, const_cast<void*>(p.W), gdim, gstr, box, est, CU_TENSOR_MAP_INTERLEAVE_NONE,
                              CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (r != CUDA_SUCCESS) { delete plan; snprintf(msg, sizeof msg, "cuTensorMapEncodeTiled(W) failed: %d", (int)r); return bad(msg); }
        plan->tmWh = plan->tmW;
        if (a.full_tiles < a.num_tiles) {
            box[1] = (cuuint32_t)(bn / 2);
            r = g_encode(&plan->tmWh, dt, 2, const_cast<void*>(p.W), gdim, gstr, box, est, CU_TENSOR_MAP_INTERLEAVE_NONE,
                         CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
            if (r != CUDA_SUCCESS) { delete plan; snprintf(msg, sizeof msg, "cuTensorMapEncodeTiled(W half) failed: %d", (int)r); return bad(msg); }
        }
    }
    // ---- row-major output through TMA stores: one box = an epilogue warp's 32 tile rows x 128 bytes ----
    a.tma_out = 0;
    a.stat_partial = nullptr;
    plan->stat_chunks = 0;
    plan->tmO = plan->tmW;
    const int cw = p.out_type == DT_F32 ? 32 : 64;
    if (p.out && bn % cw == 0 && !getenv("AVC_GEMM_NO_TMA_STORE") && ((uintptr_t)p.out % 16) == 0) {
        a.ob_t = a.TB >= 32 ? 32 : a.TB; a.ob_b = 32 / a.ob_t;
        const size_t es = dtype_size(p.out_type);
        const CUtensorMapDataType odt = p.out_type == DT_F32 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32
                                      : p.out_type == DT_BF16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT16;
        cuuint64_t gdim[3] = {(cuuint64_t)p.N, (cuuint64_t)a.gT, (cuuint64_t)a.gB};
        cuuint64_t gstr[2] = {(cuuint64_t)p.ldo * es, (cuuint64_t)a.gT * p.ldo * es};
        cuuint32_t box[3] = {(cuuint32_t)cw, (cuuint32_t)a.ob_t, (cuuint32_t)a.ob_b};
        cuuint32_t est[3] = {1, 1, 1};
        CUresult r = g_encode(&plan->tmO, odt, 3, p.out, gdim, gstr, box, est, CU_TENSOR_MAP_INTERLEAVE_NONE,
                              CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (r != CUDA_SUCCESS) { delete plan; snprintf(msg, sizeof msg, "cuTensorMapEncodeTiled(out) failed: %d", (int)r); return bad(msg); }
        a.tma_out = 1;
        // fused BatchNorm statistics: fp32 rows through the staged epilogue, full- or half-width 256/128 tiles
        if (p.stat_partial && p.out_type == DT_F32 && (bn == 256 || bn == 128) && tiles_m * 4 <= p.stat_capacity) {
            a.stat_partial = p.stat_partial;
            plan->stat_chunks = tiles_m * 4;
        }
    }
    return plan;
}

void tc_gemm_plan_destroy(TcGemmPlan* plan) { delete plan; }
int tc_gemm_plan_stat_chunks(const TcGemmPlan* plan) { return plan->stat_chunks; }

cudaError_t launch_tc_gemm(const TcGemmPlan* plan, cudaStream_t s, float* out_t, const float* res_t) {
    // the cached plan owns only workspace/weight addresses; caller-owned output pointers are per call
    TcGemmArgs args = plan->args;
    args.out_t = out_t;
    args.res_t = res_t;
    switch (plan->block_n) {
        case 256:
            if (plan->stat_chunks) tc_gemm_kernel<256, true><<<plan->grid, 192, Cfg<256>::SMEM_BYTES, s>>>(plan->tmA, plan->tmW, plan->tmO, plan->tmWh, args);
            else tc_gemm_kernel<256><<<plan->grid, 192, Cfg<256>::SMEM_BYTES, s>>>(plan->tmA, plan->tmW, plan->tmO, plan->tmWh, args);
            break;
        case 128:
            if (plan->stat_chunks) tc_gemm_kernel<128, true><<<plan->grid, 192, Cfg<128>::SMEM_BYTES, s>>>(plan->tmA, plan->tmW, plan->tmO, plan->tmWh, args);
            else tc_gemm_kernel<128><<<plan->grid, 192, Cfg<128>::SMEM_BYTES, s>>>(plan->tmA, plan->tmW, plan->tmO, plan->tmWh, args);
            break;
        case 64: tc_gemm_kernel<64><<<plan->grid, 192, Cfg<64>::SMEM_BYTES, s>>>(plan->tmA, plan->tmW, plan->tmO, plan->tmWh, args); break;
        case 32: tc_gemm_kernel<32><<<plan->grid, 192, Cfg<32>::SMEM_BYTES, s>>>(plan->tmA, plan->tmW, plan->tmO, plan->tmWh, args); break;
        default: tc_gemm_kernel<80><<<plan->grid, 192, Cfg<80>::SMEM_BYTES, s>>>(plan->tmA, plan->tmW, plan->tmO, plan->tmWh, args); break;
    }
    return cudaGetLastError();
}

}  // namespace avc
