// WaveRNN vocoder `generate` on sm_100a (reference autovc/wavernn/model.py:214-329).
//
// The per-sample loop is 12k dependent steps of  I -> GRU(512) -> GRU(512) -> fc1 -> fc2 -> fc3 -> sample : 3.2 M
// multiply-adds per step and row over 6.5 MB of bf16 weights, far more than one SM's shared memory.  So the whole GPU
// is one persistent, weight-stationary machine (cooperative launch, 128 CTAs):
//
//   CTA c owns hidden units 4c..4c+3 of EVERY layer; its slices of all weight matrices (114 KiB as ready-made 128B-
//   swizzled K-major operand blocks) stay in shared memory for the whole call.  Rows (folds of utterances) are the MMA M
//   dimension: tiles of 128 rows, up to 4 tiles per launch; accumulators live in tensor memory.
//
//   step t, four exchange phases + one local one (the operand of each phase is what ALL CTAs produced in the one before,
//   exchanged through L2 as operand tiles and brought in by 16 KiB bulk copies through a shared-memory ring):
//     A  gates of GRU 1   D[128 x 32] = [h1(t-1) | mel_t, aux1_t] . W_A^T      -> h1(t), x1 = xI + h1
//     B  gates of GRU 2   D[128 x 16] = [h2(t-1) | aux2_t | x1] . W_B^T        -> h2(t), x2 = x1 + h2
//     C  fc1 + relu       D[128 x 16] = [aux3_t | x2] . W_C^T
//     D  fc2 + relu       D[128 x 16] = [aux4_t | y1] . W_D^T
//     E  fc3 + sampling   D[128 x 32] = y2 . W_E^T, computed by EVERY CTA (30 logits cannot be split 128 ways), so the
//        new sample x_t is known locally everywhere and needs no fifth exchange.
//   Algebra done once at load: `I` is linear and feeds GRU 1 directly, so W_ih1.(W_I.[x, mel, aux1] + b_I) is folded
//   into one matrix over [mel, aux1] (K = 112) plus a rank-1 term in the scalar x that the epilogue adds on the CUDA
//   cores; the K = 512 product W_ih1.xI disappears, and h1-part and conditioning-part of phase A do not depend on x_t-1,
//   so they run while phase E of the previous step is still sampling.
//
//   Warp roles per CTA: warp 0 producer (bulk copies, grid-phase counter waits), warp 1 tcgen05.mma issue, warps 4-7
//   epilogue (thread = row; GRU state of its 4 units in registers, fp32).  A phase ends with one release-increment per
//   epilogue warp on a global counter; the producers of all CTAs acquire it before they fetch that phase's output.
//
// Every wait is bounded: a protocol bug traps (CUDA error), it cannot hang the GPU.
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>
#include <cuda.h>

#include "../../include/autovc_b200.h"
#include "common.cuh"
#include "launch.h"
#include "ptx.cuh"

namespace avc {
namespace {

// ---- model constants (reference utils/hparams.py:92-107) --------------------------------------------------------------
constexpr int V_R = 512;            // rnn_dims = fc_dims
constexpr int V_FEAT = 80;
constexpr int V_AUX = 32;           // res_out_dims / 4
constexpr int V_CD = 128;           // compute_dims = res_out_dims
constexpr int V_HOP = 275;
constexpr int V_PAD = 2;
constexpr int V_RES = 10;
constexpr int V_NCLS = 30;

// ---- decomposition ---------------------------------------------------------------------------------------------------
constexpr int VU = 4;                       // hidden units per CTA
constexpr int VCTAS = V_R / VU;             // 128 CTAs
constexpr int TILE = 128;                   // rows per tile = UMMA M
constexpr int KB = 64;                      // K elements per operand block (one 128-byte swizzle row)
constexpr int BLK = TILE * KB * 2;          // 16 KiB activation block
constexpr int NKB = V_R / KB;               // 8 blocks per 512-wide activation
constexpr int ACT_BYTES = NKB * BLK;        // 128 KiB per tile and activation
constexpr int COND_BLKS = 5;                // [mel 0..63] [mel 64..79, aux1, 0] [aux2, 0] [aux3, 0] [aux4, 0]
constexpr int MAX_TILES = 4;

// weight blocks: [rows][64] bf16, 128-byte rows, XOR-swizzled; the MMA reads N rows at 8-row-group stride 1024, so a
// matrix stored with fewer rows than N just reads on into the next block (finite weights, columns nobody looks at)
constexpr int NA = 32, ROWS_A = 24, KBA = NKB + 2;            // h1 | cond0 cond1
constexpr int NB = 16, ROWS_B = 16, KBB = NKB + 1 + NKB;      // h2 | aux2 | x1
constexpr int NC = 16, ROWS_C = 8, KBC = 1 + NKB;             // aux3 | x2       (fc2: aux4 | y1)
constexpr int NE = 32, ROWS_E = 32, KBE = NKB;
constexpr int OFF_A = 0;
constexpr int OFF_B = OFF_A + KBA * ROWS_A * 128;
constexpr int OFF_C = OFF_B + KBB * ROWS_B * 128;
constexpr int OFF_D = OFF_C + KBC * ROWS_C * 128;
constexpr int OFF_E = OFF_D + KBC * ROWS_C * 128;
constexpr int W_BYTES = OFF_E + KBE * ROWS_E * 128;            // 116,736
constexpr int NPARAM = 128;                                    // fp32 per CTA: biases and the x coefficients
constexpr int P_BA = 0, P_WXA = 20, P_BB = 40, P_BC = 56, P_BD = 60, P_BE = 64;
constexpr int STAGES = 6;
constexpr int SMEM_BYTES = 1024 + W_BYTES + STAGES * BLK + NPARAM * 4 + 256;
constexpr uint32_t TMEM_COLS = 512;
constexpr int TM_A = 0, TM_B = 32, TM_C = 48, TM_D = 64, TM_E = 96, TM_TILE = 128;   // accumulator columns of a tile
static_assert(SMEM_BYTES <= 227 * 1024, "shared memory budget");

struct VocArgs {
    int R;                          // valid rows
    int L;                          // steps
    const uint8_t* wpack;           // [VCTAS][W_BYTES]
    const float* params;            // [VCTAS][NPARAM]
    const uint8_t* cond;            // [L][tiles][COND_BLKS][BLK]
    uint8_t* h1;                    // [2][tiles][ACT_BYTES]   (double-buffered by step parity)
    uint8_t* h2;                    // [2][tiles][ACT_BYTES]
    uint8_t* x1; uint8_t* x2; uint8_t* y1; uint8_t* y2;       // [tiles][ACT_BYTES]
    unsigned int* counter;          // per-tile phase counters (128 bytes apart), zero at launch
    const float* uniforms;          // [L][R][11] or nullptr (then Philox with `seed`)
    unsigned long long seed;
    const float* forced;            // [L][R] or nullptr: fed back instead of the sample (teacher forcing)
    float* samples;                 // [R][L]
    float* logits;                  // [L][R][30] or nullptr
    uint32_t idesc32, idesc16;
    int stages;                     // ring stages in use (<= STAGES; AVC_VOC_STAGES, for experiments)
    int cs;                         // cluster size: the CTAs of a cluster fetch each activation block once and multicast it
};

__device__ __forceinline__ unsigned int ld_acquire_u32(const unsigned int* p) {
    unsigned int v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void red_release_add_u32(unsigned int* p, unsigned int v) {
    asm volatile("red.release.gpu.global.add.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ float sigmoidf_(float x) { return 1.0f / (1.0f + __expf(-x)); }

// Philox4x32-10 (Salmon et al. 2011): the stream used when the caller passes no uniforms
__device__ __forceinline__ uint4 philox(uint4 c, uint2 k) {
#pragma unroll
    for (int i = 0; i < 10; ++i) {
        const uint32_t hi0 = __umulhi(0xD2511F53u, c.x), lo0 = 0xD2511F53u * c.x;
        const uint32_t hi1 = __umulhi(0xCD9E8D57u, c.z), lo1 = 0xCD9E8D57u * c.z;
        c = make_uint4(hi1 ^ c.y ^ k.x, lo1, hi0 ^ c.w ^ k.y, lo0);
        k.x += 0x9E3779B9u; k.y += 0xBB67AE85u;
    }
    return c;
}
__device__ __forceinline__ float u01(uint32_t r) {         // U(1e-5, 1 - 1e-5), as distribution.py:106,117
    return 1e-5f + (1.0f - 2e-5f) * ((float)(r >> 8) * (1.0f / 16777216.0f));
}

template <int MT>
__global__ void __launch_bounds__(256, 1) voc_generate_kernel(const VocArgs a) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* sW = smem;
    uint8_t* sRing = smem + W_BYTES;
    float* sP = reinterpret_cast<float*>(sRing + STAGES * BLK);
    uint64_t* bars = reinterpret_cast<uint64_t*>(sP + NPARAM);
    uint64_t* full = bars;                      // [STAGES]
    uint64_t* empty = bars + STAGES;            // [STAGES]
    uint64_t* wfull = bars + 2 * STAGES;
    uint64_t* accA = wfull + 1;                 // [MAX_TILES] phase A accumulators complete
    uint64_t* accX = accA + MAX_TILES;          // [MAX_TILES] phases B..E
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(accX + MAX_TILES);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int cta = blockIdx.x;
    constexpr unsigned int ARRIVALS = 4 * VCTAS;        // one increment per epilogue warp and CTA per phase

    if (warp == 0 && lane == 0) {
        // a ring stage is refilled for the whole cluster, so it is free only when all cs CTAs have consumed it
        for (int i = 0; i < STAGES; ++i) { ptx::mbar_init(&full[i], 1); ptx::mbar_init(&empty[i], (uint32_t)a.cs); }
        ptx::mbar_init(wfull, 1);
        for (int m = 0; m < MAX_TILES; ++m) { ptx::mbar_init(&accA[m], 1); ptx::mbar_init(&accX[m], 1); }
        ptx::fence_barrier_init();
    }
    if (warp == 1) ptx::tmem_alloc<TMEM_COLS>(tmem_slot);
    if (threadIdx.x >= 128 && threadIdx.x < 128 + NPARAM) sP[threadIdx.x - 128] = a.params[(size_t)cta * NPARAM + threadIdx.x - 128];
    ptx::tc_fence_before();
    __syncthreads();
    if (a.cs > 1) ptx::cluster_sync();      // peers' barriers are initialised before anyone multicasts / commits into them
    ptx::tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    const uint32_t crank = a.cs > 1 ? ptx::cluster_ctarank() : 0u;
    const uint16_t cmask = (uint16_t)((1u << a.cs) - 1u);

    if (warp == 0) {
        // ===================== producer: the whole warp, lane i brings block i of a group =====================
        // (the issue of one bulk copy costs its thread a few hundred cycles whatever its size; eight lanes issue the eight
        // blocks of an activation side by side)
        if (lane == 0) {
            constexpr int WCH = W_BYTES / 6;                    // 19,456 B per copy
            ptx::mbar_expect_tx(wfull, (uint32_t)W_BYTES);
            for (int i = 0; i < 6; ++i) ptx::bulk_load(sW + i * WCH, a.wpack + (size_t)cta * W_BYTES + i * WCH, WCH, wfull);
        }
        int stage = 0; uint32_t phase = 0, seq = 0;
        auto load = [&](const uint8_t* src, int nblk) {
            // sub-groups of at most 4 (< STAGES) blocks with a warp sync after each: every earlier use of a stage has been
            // ISSUED before a lane waits for it to be consumed, so the parity it waits on is never two phases away
            for (int b0 = 0; b0 < nblk; b0 += min(4, a.stages - 1)) {
                const int n = min(min(4, a.stages - 1), nblk - b0);
                if (lane < n) {
                    const int pos = stage + lane;
                    const int st = pos % a.stages;
                    const uint32_t ph = phase ^ (uint32_t)(pos / a.stages);
                    ptx::mbar_wait(&empty[st], ph ^ 1);
                    ptx::mbar_expect_tx(&full[st], (uint32_t)BLK);
                    if (a.cs == 1) ptx::bulk_load(sRing + st * BLK, src + (size_t)(b0 + lane) * BLK, (uint32_t)BLK, &full[st]);
                    else if ((seq + (uint32_t)lane) % (uint32_t)a.cs == crank)      // one L2 read serves the cluster
                        ptx::bulk_load_mc(sRing + st * BLK, src + (size_t)(b0 + lane) * BLK, (uint32_t)BLK, &full[st], cmask);
                }
                seq += (uint32_t)n;
                __syncwarp();
                const int pos = stage + n;
                phase ^= (uint32_t)(pos / a.stages);
                stage = pos % a.stages;
            }
        };
        auto wait_grid = [&](int m, unsigned int k) {           // all CTAs have finished k phases of tile m
            const unsigned int target = ARRIVALS * k;
            const unsigned int* ctr = a.counter + m * 32;
            for (uint32_t spin = 0; spin < (1u << 27); ++spin)
                if (ld_acquire_u32(ctr) >= target) {
                    asm volatile("fence.proxy.async.global;" ::: "memory");   // their generic-proxy stores -> our bulk reads
                    return;
                }
            printf("autovc_b200: vocoder phase counter timed out (cta %d tile %d, want %u, have %u)\n", cta, m, target, ld_acquire_u32(ctr));
            __trap();
        };
        for (int t = 0; t < a.L; ++t) {
            const int prev = (t + 1) & 1;                                   // buffer holding h(t-1); zero-filled for t = 0
            const uint8_t* cond_t = a.cond + (size_t)t * MT * COND_BLKS * BLK;
            for (int m = 0; m < MT; ++m) {                                  // A: h1(t-1) | cond 0,1
                load(a.h1 + ((size_t)prev * MT + m) * ACT_BYTES, NKB);
                load(cond_t + (size_t)(m * COND_BLKS) * BLK, 2);
            }
            for (int m = 0; m < MT; ++m) {                                  // B, independent part: h2(t-1) | aux2
                load(a.h2 + ((size_t)prev * MT + m) * ACT_BYTES, NKB);
                load(cond_t + (size_t)(m * COND_BLKS + 2) * BLK, 1);
            }
            for (int m = 0; m < MT; ++m) { wait_grid(m, 4u * t + 1); load(a.x1 + (size_t)m * ACT_BYTES, NKB); }
            for (int m = 0; m < MT; ++m) load(cond_t + (size_t)(m * COND_BLKS + 3) * BLK, 1);       // C
            for (int m = 0; m < MT; ++m) { wait_grid(m, 4u * t + 2); load(a.x2 + (size_t)m * ACT_BYTES, NKB); }
            for (int m = 0; m < MT; ++m) load(cond_t + (size_t)(m * COND_BLKS + 4) * BLK, 1);       // D
            for (int m = 0; m < MT; ++m) { wait_grid(m, 4u * t + 3); load(a.y1 + (size_t)m * ACT_BYTES, NKB); }
            for (int m = 0; m < MT; ++m) { wait_grid(m, 4u * t + 4); load(a.y2 + (size_t)m * ACT_BYTES, NKB); }   // E
        }
    } else if (warp == 1) {
        // ===================== MMA issue =====================
        if (lane == 0) {
            ptx::mbar_wait(wfull, 0);
            int stage = 0; uint32_t phase = 0;
            // consume nblk ring blocks against weight blocks wb0.. of a matrix stored with `wstride` bytes per block
            auto mma = [&](uint32_t d_tmem, uint32_t idesc, int w_off, int wstride, int wb0, int nblk, bool fresh) {
                for (int i = 0; i < nblk; ++i) {
                    ptx::mbar_wait(&full[stage], phase);
                    ptx::tc_fence_after();
                    const uint64_t adesc = ptx::umma_desc_sw128(ptx::smem_u32(sRing + stage * BLK));
                    const uint64_t bdesc = ptx::umma_desc_sw128(ptx::smem_u32(sW + w_off + (wb0 + i) * wstride));
#pragma unroll
                    for (int k = 0; k < KB / 16; ++k)
                        ptx::umma_f16(d_tmem, adesc + (uint64_t)(k * 2), bdesc + (uint64_t)(k * 2), idesc, (fresh && i == 0 && k == 0) ? 0u : 1u);
                    if (a.cs == 1) ptx::umma_commit(&empty[stage]);
                    else ptx::umma_commit_mc(&empty[stage], cmask);
                    if (++stage == a.stages) { stage = 0; phase ^= 1; }
                }
            };
            for (int t = 0; t < a.L; ++t) {
                for (int m = 0; m < MT; ++m) {
                    mma(tmem_base + m * TM_TILE + TM_A, a.idesc32, OFF_A, ROWS_A * 128, 0, KBA, true);
                    ptx::umma_commit(&accA[m]);
                }
                for (int m = 0; m < MT; ++m) mma(tmem_base + m * TM_TILE + TM_B, a.idesc16, OFF_B, ROWS_B * 128, 0, NKB + 1, true);
                for (int m = 0; m < MT; ++m) {
                    mma(tmem_base + m * TM_TILE + TM_B, a.idesc16, OFF_B, ROWS_B * 128, NKB + 1, NKB, false);
                    ptx::umma_commit(&accX[m]);
                }
                for (int m = 0; m < MT; ++m) mma(tmem_base + m * TM_TILE + TM_C, a.idesc16, OFF_C, ROWS_C * 128, 0, 1, true);
                for (int m = 0; m < MT; ++m) {
                    mma(tmem_base + m * TM_TILE + TM_C, a.idesc16, OFF_C, ROWS_C * 128, 1, NKB, false);
                    ptx::umma_commit(&accX[m]);
                }
                for (int m = 0; m < MT; ++m) mma(tmem_base + m * TM_TILE + TM_D, a.idesc16, OFF_D, ROWS_C * 128, 0, 1, true);
                for (int m = 0; m < MT; ++m) {
                    mma(tmem_base + m * TM_TILE + TM_D, a.idesc16, OFF_D, ROWS_C * 128, 1, NKB, false);
                    ptx::umma_commit(&accX[m]);
                }
                for (int m = 0; m < MT; ++m) {
                    mma(tmem_base + m * TM_TILE + TM_E, a.idesc32, OFF_E, ROWS_E * 128, 0, KBE, true);
                    ptx::umma_commit(&accX[m]);
                }
            }
        }
    } else if (warp >= 4) {
        // ===================== epilogue: thread = row =====================
        const int q = warp & 3;
        const int r = q * 32 + lane;
        const uint32_t tbase = tmem_base + ((uint32_t)(q * 32) << 16);
        // this CTA's 8 bytes of row r inside an activation tile: K block cta/16, 16-byte chunk (cta%16)/2 swizzled by the row
        const size_t act_off = (size_t)(cta >> 4) * BLK + (size_t)r * 128 + (size_t)(((((cta & 15) >> 1) ^ (r & 7)) << 4) + ((cta & 1) << 3));
        float h1[MT][VU], h2[MT][VU], x1[MT][VU], x[MT];
#pragma unroll
        for (int m = 0; m < MT; ++m) {
            x[m] = 0.0f;
#pragma unroll
            for (int j = 0; j < VU; ++j) h1[m][j] = h2[m][j] = x1[m][j] = 0.0f;
        }
        uint32_t parA = 0, parX = 0;
        auto pack4 = [](const float (&v)[VU]) {
            __nv_bfloat162 lo = __floats2bfloat162_rn(v[0], v[1]), hi = __floats2bfloat162_rn(v[2], v[3]);
            return make_uint2(*reinterpret_cast<uint32_t*>(&lo), *reinterpret_cast<uint32_t*>(&hi));
        };
        auto publish = [&](int m) {                 // this warp's rows of tile m are stored: one arrival on the tile's counter
            asm volatile("fence.proxy.async.global;" ::: "memory");
            __syncwarp();
            if (lane == 0) red_release_add_u32(a.counter + m * 32, 1u);
        };
        for (int t = 0; t < a.L; ++t) {
            const int cur = t & 1;
            // the step's random numbers (or the forced sample) do not depend on anything: fetch them first
            float u[MT][11], fx[MT];
#pragma unroll
            for (int m = 0; m < MT; ++m) {
                const int row = m * TILE + r;
                const bool valid = row < a.R;
                fx[m] = (a.forced && valid) ? __ldg(a.forced + (size_t)t * a.R + row) : 0.0f;
                if (a.uniforms) {
#pragma unroll
                    for (int k = 0; k < 11; ++k) u[m][k] = valid ? __ldg(a.uniforms + ((size_t)t * a.R + row) * 11 + k) : 0.5f;
                } else {
                    const uint2 key = make_uint2((uint32_t)a.seed, (uint32_t)(a.seed >> 32));
#pragma unroll
                    for (int g = 0; g < 3; ++g) {
                        const uint4 rr = philox(make_uint4((uint32_t)t, (uint32_t)row, (uint32_t)g, 0x57a7e5u), key);
                        u[m][4 * g] = u01(rr.x); u[m][4 * g + 1] = u01(rr.y); u[m][4 * g + 2] = u01(rr.z);
                        if (g < 2) u[m][4 * g + 3] = u01(rr.w);
                    }
                }
            }
            // ---- A: GRU 1 ----
#pragma unroll
            for (int m = 0; m < MT; ++m) {
                ptx::mbar_wait(&accA[m], parA);
                ptx::tc_fence_after();
                uint32_t d0[16], d1[16];
                ptx::tmem_ld16(tbase + m * TM_TILE + TM_A, d0);
                ptx::tmem_ld16(tbase + m * TM_TILE + TM_A + 16, d1);
                ptx::tmem_ld_wait();
                float hv[VU];
#pragma unroll
                for (int j = 0; j < VU; ++j) {
                    const float pr = __uint_as_float(d0[j]) + x[m] * sP[P_WXA + j] + sP[P_BA + j];
                    const float pz = __uint_as_float(d0[4 + j]) + x[m] * sP[P_WXA + 4 + j] + sP[P_BA + 4 + j];
                    const float ph = __uint_as_float(d0[8 + j]) + sP[P_BA + 8 + j];
                    const float pi = __uint_as_float(d0[12 + j]) + x[m] * sP[P_WXA + 12 + j] + sP[P_BA + 12 + j];
                    const float xi = __uint_as_float(d1[j]) + x[m] * sP[P_WXA + 16 + j] + sP[P_BA + 16 + j];
                    const float rg = sigmoidf_(pr), zg = sigmoidf_(pz);
                    const float ng = tanhf(pi + rg * ph);
                    h1[m][j] = (1.0f - zg) * ng + zg * h1[m][j];
                    x1[m][j] = xi + h1[m][j];
                    hv[j] = h1[m][j];
                }
                *reinterpret_cast<uint2*>(a.h1 + ((size_t)cur * MT + m) * ACT_BYTES + act_off) = pack4(hv);
                *reinterpret_cast<uint2*>(a.x1 + (size_t)m * ACT_BYTES + act_off) = pack4(x1[m]);
                publish(m);
            }
            parA ^= 1;
            // ---- B: GRU 2 ----
#pragma unroll
            for (int m = 0; m < MT; ++m) {
                ptx::mbar_wait(&accX[m], parX);
                ptx::tc_fence_after();
                uint32_t d0[16];
                ptx::tmem_ld16(tbase + m * TM_TILE + TM_B, d0);
                ptx::tmem_ld_wait();
                float hv[VU];
#pragma unroll
                for (int j = 0; j < VU; ++j) {
                    const float rg = sigmoidf_(__uint_as_float(d0[j]) + sP[P_BB + j]);
                    const float zg = sigmoidf_(__uint_as_float(d0[4 + j]) + sP[P_BB + 4 + j]);
                    const float ng = tanhf(__uint_as_float(d0[12 + j]) + sP[P_BB + 12 + j] + rg * (__uint_as_float(d0[8 + j]) + sP[P_BB + 8 + j]));
                    h2[m][j] = (1.0f - zg) * ng + zg * h2[m][j];
                    x1[m][j] += h2[m][j];                   // x2 (x1 is not needed again)
                    hv[j] = h2[m][j];
                }
                *reinterpret_cast<uint2*>(a.h2 + ((size_t)cur * MT + m) * ACT_BYTES + act_off) = pack4(hv);
                *reinterpret_cast<uint2*>(a.x2 + (size_t)m * ACT_BYTES + act_off) = pack4(x1[m]);
                publish(m);
            }
            parX ^= 1;
            // ---- C, D: fc1, fc2 ----
#pragma unroll
            for (int ph = 0; ph < 2; ++ph) {
#pragma unroll
                for (int m = 0; m < MT; ++m) {
                    ptx::mbar_wait(&accX[m], parX);
                    ptx::tc_fence_after();
                    uint32_t d0[16];
                    ptx::tmem_ld16(tbase + m * TM_TILE + (ph ? TM_D : TM_C), d0);
                    ptx::tmem_ld_wait();
                    float y[VU];
#pragma unroll
                    for (int j = 0; j < VU; ++j) y[j] = fmaxf(__uint_as_float(d0[j]) + sP[(ph ? P_BD : P_BC) + j], 0.0f);
                    *reinterpret_cast<uint2*>((ph ? a.y2 : a.y1) + (size_t)m * ACT_BYTES + act_off) = pack4(y);
                    publish(m);
                }
                parX ^= 1;
            }
            // ---- E: fc3 + mixture-of-logistics sample (distribution.py:87-123), identical in every CTA ----
#pragma unroll
            for (int m = 0; m < MT; ++m) {
                ptx::mbar_wait(&accX[m], parX);
                ptx::tc_fence_after();
                uint32_t d0[16], d1[16];
                ptx::tmem_ld16(tbase + m * TM_TILE + TM_E, d0);
                ptx::tmem_ld16(tbase + m * TM_TILE + TM_E + 16, d1);
                ptx::tmem_ld_wait();
                float lg[32];
#pragma unroll
                for (int k = 0; k < 16; ++k) { lg[k] = __uint_as_float(d0[k]) + sP[P_BE + k]; lg[16 + k] = __uint_as_float(d1[k]) + sP[P_BE + 16 + k]; }
                int best = 0; float bv = -INFINITY, mean = 0.0f, ls = 0.0f;
#pragma unroll
                for (int k = 0; k < 10; ++k) {
                    const float v = lg[k] - logf(-logf(u[m][k]));
                    if (v > bv) { bv = v; best = k; mean = lg[10 + k]; ls = lg[20 + k]; }
                }
                (void)best;
                ls = fmaxf(ls, -32.236191301916641f);                       // log(1e-14)
                float s = mean + expf(ls) * (logf(u[m][10]) - logf(1.0f - u[m][10]));
                s = fminf(fmaxf(s, -1.0f), 1.0f);
                const int row = m * TILE + r;
                if (cta == 0 && row < a.R) {
                    a.samples[(size_t)row * a.L + t] = s;
                    if (a.logits) {
#pragma unroll
                        for (int k = 0; k < V_NCLS; ++k) a.logits[((size_t)t * a.R + row) * V_NCLS + k] = lg[k];
                    }
                }
                x[m] = a.forced ? fx[m] : s;
            }
            parX ^= 1;
        }
    }

    ptx::tc_fence_before();
    __syncthreads();
    if (a.cs > 1) ptx::cluster_sync();      // no CTA may exit while peers can still multicast / commit into it
    if (warp == 1) {
        ptx::tc_fence_after();
        ptx::tmem_dealloc<TMEM_COLS>(tmem_base);
    }
}

// ---- conditioning network (UpsampleNetwork, model.py:66-92), fp32 on the CUDA cores: 0.02 % of the work ---------------------
// MelResNet: conv_in (k=5, BatchNorm folded) -> relu -> 10 residual blocks of 1x1 convolutions -> conv_out.  One CTA per
// (utterance, 16 frames); thread = output channel; activations in shared memory; weights pre-transposed [ci][co].
constexpr int RF = 16;
struct ResnetArgs {
    const float* mel;      // [U][80][T]   (un-padded; the `pad` zero frames either side are applied by index)
    float* aux;            // [U][T][128]
    const float* w_in;     // [80*5][128]
    const float* b_in;     // [128]
    const float* w_res;    // [10][2][128 ci][128 co]
    const float* b_res;    // [10][2][128]
    const float* w_out;    // [128][128]
    const float* b_out;    // [128]
    int U, T;
};
__global__ void __launch_bounds__(V_CD) voc_resnet_kernel(const ResnetArgs a) {
    __shared__ float sIn[V_FEAT][RF + 2 * V_PAD];
    __shared__ float sX[V_CD][RF + 1];
    __shared__ float sY[V_CD][RF + 1];
    const int u = blockIdx.y, f0 = blockIdx.x * RF, co = threadIdx.x;
    for (int i = threadIdx.x; i < V_FEAT * (RF + 2 * V_PAD); i += V_CD) {
        const int c = i / (RF + 2 * V_PAD), k = i % (RF + 2 * V_PAD);
        const int f = f0 + k - V_PAD;
        sIn[c][k] = (f >= 0 && f < a.T) ? a.mel[((size_t)u * V_FEAT + c) * a.T + f] : 0.0f;
    }
    __syncthreads();
    float acc[RF];
#pragma unroll
    for (int f = 0; f < RF; ++f) acc[f] = a.b_in[co];
    for (int c = 0; c < V_FEAT; ++c)
#pragma unroll
        for (int j = 0; j < 2 * V_PAD + 1; ++j) {
            const float w = a.w_in[(size_t)(c * (2 * V_PAD + 1) + j) * V_CD + co];
#pragma unroll
            for (int f = 0; f < RF; ++f) acc[f] = fmaf(w, sIn[c][f + j], acc[f]);
        }
#pragma unroll
    for (int f = 0; f < RF; ++f) sX[co][f] = fmaxf(acc[f], 0.0f);
    __syncthreads();
    for (int l = 0; l < V_RES; ++l) {
        const float* w1 = a.w_res + (size_t)(l * 2) * V_CD * V_CD;
        const float* w2 = w1 + V_CD * V_CD;
#pragma unroll
        for (int f = 0; f < RF; ++f) acc[f] = a.b_res[(l * 2) * V_CD + co];
        for (int c = 0; c < V_CD; ++c) {
            const float w = w1[(size_t)c * V_CD + co];
#pragma unroll
            for (int f = 0; f < RF; ++f) acc[f] = fmaf(w, sX[c][f], acc[f]);
        }
#pragma unroll
        for (int f = 0; f < RF; ++f) sY[co][f] = fmaxf(acc[f], 0.0f);
        __syncthreads();
#pragma unroll
        for (int f = 0; f < RF; ++f) acc[f] = a.b_res[(l * 2 + 1) * V_CD + co] + sX[co][f];
        for (int c = 0; c < V_CD; ++c) {
            const float w = w2[(size_t)c * V_CD + co];
#pragma unroll
            for (int f = 0; f < RF; ++f) acc[f] = fmaf(w, sY[c][f], acc[f]);
        }
        __syncthreads();                     // everyone has read sX
#pragma unroll
        for (int f = 0; f < RF; ++f) sX[co][f] = acc[f];
        __syncthreads();
    }
#pragma unroll
    for (int f = 0; f < RF; ++f) acc[f] = a.b_out[co];
    for (int c = 0; c < V_CD; ++c) {
        const float w = a.w_out[(size_t)c * V_CD + co];
#pragma unroll
        for (int f = 0; f < RF; ++f) acc[f] = fmaf(w, sX[c][f], acc[f]);
    }
    for (int f = 0; f < RF; ++f)
        if (f0 + f < a.T) a.aux[((size_t)u * a.T + f0 + f) * V_CD + co] = acc[f];
}

// One Stretch2d(s, 1) + Conv2d(1, 1, (1, 2s+1), padding (0, s)) layer (model.py:77-84) along time, shared by the 80 mel
// channels: out[p] = sum_j w[j] * stretched[p + lo + j - s], stretched[q] = in[q / s] for 0 <= q < Lin*s, else 0.
// in_pad > 0: `in` is the un-padded mel and in_pad zero frames sit either side (pad_tensor, model.py:252).
// transposed: out is [U][Lout][80] (the layout the loop reads), else [U][80][Lout].
__global__ void voc_stretch_conv_kernel(const float* __restrict__ in, float* __restrict__ out, const float* __restrict__ w, int s, int U,
                                        int Lin_data, int in_pad, long long lo, long long Lout, int transposed) {
    const long long n = (long long)U * V_FEAT * Lout;
    const long long Lin = Lin_data + 2 * in_pad;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        int c; long long p; int u;
        if (transposed) { c = (int)(i % V_FEAT); p = (i / V_FEAT) % Lout; u = (int)(i / (V_FEAT * Lout)); }
        else { p = i % Lout; c = (int)((i / Lout) % V_FEAT); u = (int)(i / (Lout * V_FEAT)); }
        const float* row = in + ((size_t)u * V_FEAT + c) * Lin_data;
        float acc = 0.0f;
        for (int j = 0; j <= 2 * s; ++j) {
            const long long qq = p + lo + j - s;
            if (qq < 0 || qq >= Lin * s) continue;
            const long long f = qq / s - in_pad;
            if (f >= 0 && f < Lin_data) acc = fmaf(w[j], row[f], acc);
        }
        out[i] = acc;
    }
}

// Rows of the loop = windows (utterance, start) of L steps over the conditioning, zero past its end (pad_tensor 'after'
// + fold_with_overlap, model.py:358-401).  Writes the per-step operand blocks: [L][tiles][5][128 rows][64] bf16, swizzled.
__global__ void voc_pack_cond_kernel(const float* __restrict__ m_up, const float* __restrict__ aux, const int* __restrict__ row_utt,
                                     const int* __restrict__ row_start, uint8_t* __restrict__ cond, int R, int tiles, int L, long long Ltot,
                                     int T) {
    // one thread per 16-byte chunk: 5 blocks x 8 chunks per row and step
    const long long n = (long long)L * tiles * TILE * (COND_BLKS * 8);
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        const int ch = (int)(i % 8);
        const int r = (int)((i / 8) % TILE);
        const int blk = (int)((i / (8 * TILE)) % COND_BLKS);
        const int m = (int)((i / (8 * TILE * COND_BLKS)) % tiles);
        const long long t = i / ((long long)8 * TILE * COND_BLKS * tiles);
        const int row = m * TILE + r;
        float v[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) v[k] = 0.0f;
        if (row < R) {
            const long long p = (long long)row_start[row] + t;
            const int u = row_utt[row];
            if (p < Ltot) {
                const float* mp = m_up + ((size_t)u * Ltot + p) * V_FEAT;
                const float* ap = aux + ((size_t)u * T + p / V_HOP) * V_CD;
                const int k0 = ch * 8;
                if (blk == 0) {
#pragma unroll
                    for (int k = 0; k < 8; ++k) v[k] = mp[k0 + k];
                } else if (blk == 1) {
                    if (k0 < 16) {
#pragma unroll
                        for (int k = 0; k < 8; ++k) v[k] = mp[64 + k0 + k];
                    } else if (k0 < 48) {
#pragma unroll
                        for (int k = 0; k < 8; ++k) v[k] = ap[k0 - 16 + k];
                    }
                } else if (k0 < 32) {
#pragma unroll
                    for (int k = 0; k < 8; ++k) v[k] = ap[(blk - 1) * V_AUX + k0 + k];
                }
            }
        }
        __nv_bfloat162 o[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) o[k] = __floats2bfloat162_rn(v[2 * k], v[2 * k + 1]);
        uint8_t* dst = cond + (((size_t)t * tiles + m) * COND_BLKS + blk) * BLK + (size_t)r * 128 + (size_t)((ch ^ (r & 7)) << 4);
        *reinterpret_cast<uint4*>(dst) = *reinterpret_cast<uint4*>(o);
    }
}

// ---- host: weight folding and packing -----------------------------------------------------------------------------------
uint16_t f2bf(float f) {
    uint32_t u; memcpy(&u, &f, 4);
    if ((u & 0x7fffffffu) > 0x7f800000u) return (uint16_t)((u >> 16) | 0x40);
    u += 0x7fffu + ((u >> 16) & 1u);
    return (uint16_t)(u >> 16);
}
// element (row, k) of weight block `kb` of a matrix stored with `rows` rows per block
inline void put(std::vector<uint16_t>& img, int off_bytes, int rows, int kb, int row, int k, double v) {
    const size_t byte = (size_t)off_bytes + (size_t)kb * rows * 128 + (size_t)row * 128 + (size_t)((((k >> 3) ^ (row & 7)) << 4) + ((k & 7) << 1));
    img[byte / 2] = f2bf((float)v);
}

thread_local std::string g_voc_err;

}  // namespace
}  // namespace avc

using namespace avc;

struct avc_voc_ctx {
    int device = 0;
    int num_sms = 0;
    bool loaded = false;
    uint8_t* d_wpack = nullptr; float* d_params = nullptr;
    float *d_w_in = nullptr, *d_b_in = nullptr, *d_w_res = nullptr, *d_b_res = nullptr, *d_w_out = nullptr, *d_b_out = nullptr, *d_up = nullptr;
    int cs = 1;                 // cluster size of the loop kernel (multicast of the activation blocks)
    int64_t launches = 0;
    std::string err;
};

namespace {
int voc_fail(avc_voc_ctx* h, int code, const char* fmt, ...) {
    char buf[512];
    va_list ap; va_start(ap, fmt); vsnprintf(buf, sizeof buf, fmt, ap); va_end(ap);
    if (h) h->err = buf; else g_voc_err = buf;
    return code;
}
template <class T> cudaError_t voc_upload(T** dst, const std::vector<T>& v) {
    if (*dst) { cudaFree(*dst); *dst = nullptr; }
    cudaError_t e = cudaMalloc(dst, v.size() * sizeof(T) + 16);
    if (e != cudaSuccess) return e;
    return cudaMemcpy(*dst, v.data(), v.size() * sizeof(T), cudaMemcpyHostToDevice);
}
struct Layout { size_t cond, h1, h2, x1, x2, y1, y2, counter, total; };
Layout voc_layout(int tiles, int L) {
    Layout o{};
    size_t p = 0;
    auto take = [&](size_t bytes) { size_t at = p; p += (bytes + 1023) & ~size_t(1023); return at; };
    o.cond = take((size_t)L * tiles * COND_BLKS * BLK);
    o.h1 = take((size_t)2 * tiles * ACT_BYTES);
    o.h2 = take((size_t)2 * tiles * ACT_BYTES);
    o.x1 = take((size_t)tiles * ACT_BYTES);
    o.x2 = take((size_t)tiles * ACT_BYTES);
    o.y1 = take((size_t)tiles * ACT_BYTES);
    o.y2 = take((size_t)tiles * ACT_BYTES);
    o.counter = take(1024);
    o.total = p;
    return o;
}
int tiles_for(int rows) { const int t = ceil_div(rows, TILE); return t <= 1 ? 1 : t <= 2 ? 2 : 4; }
}  // namespace

extern "C" {

int avc_voc_create(avc_voc_handle* out, int device) {
    if (!out) return voc_fail(nullptr, AVC_ERR_INVALID_ARG, "null out");
    avc_voc_ctx* h = new avc_voc_ctx();
    h->device = device;
    if (device >= 0) {
        cudaDeviceProp prop;
        if (cudaGetDeviceProperties(&prop, device) != cudaSuccess || cudaSetDevice(device) != cudaSuccess) {
            delete h;
            return voc_fail(nullptr, AVC_ERR_CUDA, "no CUDA device %d", device);
        }
        if (prop.major != 10) {
            delete h;
            return voc_fail(nullptr, AVC_ERR_CUDA, "device %d is sm_%d%d; this library is built for sm_100a only", device, prop.major, prop.minor);
        }
        h->num_sms = prop.multiProcessorCount;
        if (h->num_sms < VCTAS) {
            delete h;
            return voc_fail(nullptr, AVC_ERR_CUDA, "the vocoder loop needs %d co-resident CTAs, the device has %d SMs", VCTAS, h->num_sms);
        }
        cudaFuncSetAttribute(voc_generate_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES);
        cudaFuncSetAttribute(voc_generate_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES);
        cudaFuncSetAttribute(voc_generate_kernel<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES);
        // largest cluster (<= 8, or AVC_VOC_CLUSTER) for which all 128 CTAs are still co-resident
        const char* env = getenv("AVC_VOC_CLUSTER");
        for (int c = env ? atoi(env) : 8; c >= 2; c >>= 1) {
            if (c > 8 || (c & (c - 1))) continue;
            cudaLaunchConfig_t cfg{};
            cfg.gridDim = dim3(VCTAS); cfg.blockDim = dim3(256); cfg.dynamicSmemBytes = SMEM_BYTES;
            cudaLaunchAttribute at[1];
            at[0].id = cudaLaunchAttributeClusterDimension;
            at[0].val.clusterDim.x = (unsigned)c; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
            cfg.attrs = at; cfg.numAttrs = 1;
            int ncl = 0;
            if (cudaOccupancyMaxActiveClusters(&ncl, voc_generate_kernel<1>, &cfg) == cudaSuccess && ncl * c >= VCTAS) { h->cs = c; break; }
        }
        cudaGetLastError();
    }
    *out = h;
    return AVC_OK;
}

int avc_voc_load_weights(avc_voc_handle h, const avc_weight_desc* table, int n) {
    if (!h || !table) return voc_fail(h, AVC_ERR_INVALID_ARG, "null handle/table");
    auto find = [&](const std::string& name, int64_t numel) -> const float* {
        for (int i = 0; i < n; ++i)
            if (name == table[i].name) {
                if (table[i].numel != numel) { voc_fail(h, AVC_ERR_MISSING_WEIGHT, "%s: %lld elements, expected %lld", name.c_str(), (long long)table[i].numel, (long long)numel); return nullptr; }
                return table[i].data;
            }
        voc_fail(h, AVC_ERR_MISSING_WEIGHT, "missing state-dict entry %s", name.c_str());
        return nullptr;
    };
#define GET(var, name, numel) const float* var = find(name, numel); if (!var) return AVC_ERR_MISSING_WEIGHT;
    const int KI = 1 + V_FEAT + V_AUX;      // 113
    GET(WI, "I.weight", (int64_t)V_R * KI) GET(bI, "I.bias", V_R)
    GET(Wi1, "rnn1.weight_ih_l0", (int64_t)3 * V_R * V_R) GET(Wh1, "rnn1.weight_hh_l0", (int64_t)3 * V_R * V_R)
    GET(bi1, "rnn1.bias_ih_l0", 3 * V_R) GET(bh1, "rnn1.bias_hh_l0", 3 * V_R)
    GET(Wi2, "rnn2.weight_ih_l0", (int64_t)3 * V_R * (V_R + V_AUX)) GET(Wh2, "rnn2.weight_hh_l0", (int64_t)3 * V_R * V_R)
    GET(bi2, "rnn2.bias_ih_l0", 3 * V_R) GET(bh2, "rnn2.bias_hh_l0", 3 * V_R)
    GET(W1, "fc1.weight", (int64_t)V_R * (V_R + V_AUX)) GET(b1, "fc1.bias", V_R)
    GET(W2, "fc2.weight", (int64_t)V_R * (V_R + V_AUX)) GET(b2, "fc2.bias", V_R)
    GET(W3, "fc3.weight", (int64_t)V_NCLS * V_R) GET(b3, "fc3.bias", V_NCLS)
    GET(cin, "upsample.resnet.conv_in.weight", (int64_t)V_CD * V_FEAT * 5)
    GET(cout, "upsample.resnet.conv_out.weight", (int64_t)V_CD * V_CD) GET(cout_b, "upsample.resnet.conv_out.bias", V_CD)
    GET(up1, "upsample.up_layers.1.weight", 11) GET(up3, "upsample.up_layers.3.weight", 11) GET(up5, "upsample.up_layers.5.weight", 23)

    // ---- the loop's matrices ----
    // F[g][k] = sum_i W_ih1[g][i] * W_I[i][k]  for the 3*512 gate rows g and the 113 inputs k of I; fb[g] = W_ih1[g] . b_I
    std::vector<double> F((size_t)3 * V_R * KI), fb((size_t)3 * V_R);
    for (int g = 0; g < 3 * V_R; ++g) {
        double* f = &F[(size_t)g * KI];
        double bsum = 0.0;
        for (int k = 0; k < KI; ++k) f[k] = 0.0;
        for (int i = 0; i < V_R; ++i) {
            const double w = Wi1[(size_t)g * V_R + i];
            const float* wi = WI + (size_t)i * KI;
            for (int k = 0; k < KI; ++k) f[k] += w * wi[k];
            bsum += w * bI[i];
        }
        fb[g] = bsum;
    }
    std::vector<uint16_t> img((size_t)VCTAS * W_BYTES / 2, 0);
    std::vector<float> params((size_t)VCTAS * NPARAM, 0.0f);
    // conditioning columns of phase A: k = 0..79 mel, 80..111 aux1  ->  block NKB (k < 64) or NKB+1 (k - 64)
    auto putA_cond = [&](int base, int row, int k, double v) { put(img, base + OFF_A, ROWS_A, NKB + (k >> 6), row, k & 63, v); };
    for (int c = 0; c < VCTAS; ++c) {
        const int base = c * W_BYTES;
        float* P = &params[(size_t)c * NPARAM];
        for (int q = 0; q < VU; ++q) {
            const int j = c * VU + q;
            const int gr = j, gz = V_R + j, gn = 2 * V_R + j;
            // A rows: q r | 4+q z | 8+q n(hidden part) | 12+q n(input part) | 16+q xI
            for (int k = 0; k < V_R; ++k) {
                put(img, base + OFF_A, ROWS_A, k >> 6, q, k & 63, Wh1[(size_t)gr * V_R + k]);
                put(img, base + OFF_A, ROWS_A, k >> 6, 4 + q, k & 63, Wh1[(size_t)gz * V_R + k]);
                put(img, base + OFF_A, ROWS_A, k >> 6, 8 + q, k & 63, Wh1[(size_t)gn * V_R + k]);
            }
            for (int k = 0; k < V_FEAT + V_AUX; ++k) {
                putA_cond(base, q, k, F[(size_t)gr * KI + 1 + k]);
                putA_cond(base, 4 + q, k, F[(size_t)gz * KI + 1 + k]);
                putA_cond(base, 12 + q, k, F[(size_t)gn * KI + 1 + k]);
                putA_cond(base, 16 + q, k, WI[(size_t)j * KI + 1 + k]);
            }
            P[P_WXA + q] = (float)F[(size_t)gr * KI]; P[P_WXA + 4 + q] = (float)F[(size_t)gz * KI]; P[P_WXA + 8 + q] = 0.0f;
            P[P_WXA + 12 + q] = (float)F[(size_t)gn * KI]; P[P_WXA + 16 + q] = WI[(size_t)j * KI];
            P[P_BA + q] = (float)((double)bi1[gr] + bh1[gr] + fb[gr]); P[P_BA + 4 + q] = (float)((double)bi1[gz] + bh1[gz] + fb[gz]);
            P[P_BA + 8 + q] = bh1[gn]; P[P_BA + 12 + q] = (float)((double)bi1[gn] + fb[gn]); P[P_BA + 16 + q] = bI[j];
            // B rows: q r | 4+q z | 8+q n(hidden) | 12+q n(input);  K blocks: 0..7 h2, 8 aux2, 9..16 x1
            const int K2 = V_R + V_AUX;
            for (int k = 0; k < V_R; ++k) {
                put(img, base + OFF_B, ROWS_B, k >> 6, q, k & 63, Wh2[(size_t)gr * V_R + k]);
                put(img, base + OFF_B, ROWS_B, k >> 6, 4 + q, k & 63, Wh2[(size_t)gz * V_R + k]);
                put(img, base + OFF_B, ROWS_B, k >> 6, 8 + q, k & 63, Wh2[(size_t)gn * V_R + k]);
                put(img, base + OFF_B, ROWS_B, NKB + 1 + (k >> 6), q, k & 63, Wi2[(size_t)gr * K2 + k]);
                put(img, base + OFF_B, ROWS_B, NKB + 1 + (k >> 6), 4 + q, k & 63, Wi2[(size_t)gz * K2 + k]);
                put(img, base + OFF_B, ROWS_B, NKB + 1 + (k >> 6), 12 + q, k & 63, Wi2[(size_t)gn * K2 + k]);
            }
            for (int k = 0; k < V_AUX; ++k) {
                put(img, base + OFF_B, ROWS_B, NKB, q, k, Wi2[(size_t)gr * K2 + V_R + k]);
                put(img, base + OFF_B, ROWS_B, NKB, 4 + q, k, Wi2[(size_t)gz * K2 + V_R + k]);
                put(img, base + OFF_B, ROWS_B, NKB, 12 + q, k, Wi2[(size_t)gn * K2 + V_R + k]);
            }
            P[P_BB + q] = (float)((double)bi2[gr] + bh2[gr]); P[P_BB + 4 + q] = (float)((double)bi2[gz] + bh2[gz]);
            P[P_BB + 8 + q] = bh2[gn]; P[P_BB + 12 + q] = bi2[gn];
            // C / D rows q: K blocks 0 aux, 1..8 activation
            for (int k = 0; k < V_R; ++k) {
                put(img, base + OFF_C, ROWS_C, 1 + (k >> 6), q, k & 63, W1[(size_t)j * K2 + k]);
                put(img, base + OFF_D, ROWS_C, 1 + (k >> 6), q, k & 63, W2[(size_t)j * K2 + k]);
            }
            for (int k = 0; k < V_AUX; ++k) {
                put(img, base + OFF_C, ROWS_C, 0, q, k, W1[(size_t)j * K2 + V_R + k]);
                put(img, base + OFF_D, ROWS_C, 0, q, k, W2[(size_t)j * K2 + V_R + k]);
            }
            P[P_BC + q] = b1[j]; P[P_BD + q] = b2[j];
        }
        for (int row = 0; row < V_NCLS; ++row) {
            for (int k = 0; k < V_R; ++k) put(img, base + OFF_E, ROWS_E, k >> 6, row, k & 63, W3[(size_t)row * V_R + k]);
            P[P_BE + row] = b3[row];
        }
    }

    // ---- conditioning network: BatchNorm (eval) folded, weights transposed to [ci][co] ----
    auto bn_fold = [&](const std::string& p, std::vector<double>& scale, std::vector<double>& shift) -> bool {
        const float* g = find(p + ".weight", V_CD); const float* b = find(p + ".bias", V_CD);
        const float* m = find(p + ".running_mean", V_CD); const float* v = find(p + ".running_var", V_CD);
        if (!g || !b || !m || !v) return false;
        scale.resize(V_CD); shift.resize(V_CD);
        for (int c = 0; c < V_CD; ++c) { scale[c] = g[c] / std::sqrt((double)v[c] + 1e-5); shift[c] = b[c] - m[c] * scale[c]; }
        return true;
    };
    std::vector<double> sc, sh;
    std::vector<float> w_in((size_t)V_FEAT * 5 * V_CD), b_in(V_CD), w_res((size_t)V_RES * 2 * V_CD * V_CD), b_res((size_t)V_RES * 2 * V_CD),
        w_out((size_t)V_CD * V_CD), b_out(cout_b, cout_b + V_CD), up(11 + 11 + 23);
    if (!bn_fold("upsample.resnet.batch_norm", sc, sh)) return AVC_ERR_MISSING_WEIGHT;
    for (int co = 0; co < V_CD; ++co) {
        for (int ci = 0; ci < V_FEAT; ++ci)
            for (int j = 0; j < 5; ++j) w_in[(size_t)(ci * 5 + j) * V_CD + co] = (float)(cin[((size_t)co * V_FEAT + ci) * 5 + j] * sc[co]);
        b_in[co] = (float)sh[co];
    }
    for (int l = 0; l < V_RES; ++l)
        for (int s = 0; s < 2; ++s) {
            const std::string p = "upsample.resnet.layers." + std::to_string(l);
            const float* w = find(p + (s ? ".conv2.weight" : ".conv1.weight"), (int64_t)V_CD * V_CD);
            if (!w || !bn_fold(p + (s ? ".batch_norm2" : ".batch_norm1"), sc, sh)) return AVC_ERR_MISSING_WEIGHT;
            for (int co = 0; co < V_CD; ++co) {
                for (int ci = 0; ci < V_CD; ++ci) w_res[((size_t)(l * 2 + s) * V_CD + ci) * V_CD + co] = (float)(w[(size_t)co * V_CD + ci] * sc[co]);
                b_res[(size_t)(l * 2 + s) * V_CD + co] = (float)sh[co];
            }
        }
    for (int co = 0; co < V_CD; ++co)
        for (int ci = 0; ci < V_CD; ++ci) w_out[(size_t)ci * V_CD + co] = cout[(size_t)co * V_CD + ci];
    memcpy(&up[0], up1, 11 * 4); memcpy(&up[11], up3, 11 * 4); memcpy(&up[22], up5, 23 * 4);
#undef GET
    if (h->device >= 0) {
        cudaSetDevice(h->device);
        std::vector<uint8_t> bytes((const uint8_t*)img.data(), (const uint8_t*)img.data() + img.size() * 2);
        if (voc_upload(&h->d_wpack, bytes) != cudaSuccess || voc_upload(&h->d_params, params) != cudaSuccess ||
            voc_upload(&h->d_w_in, w_in) != cudaSuccess || voc_upload(&h->d_b_in, b_in) != cudaSuccess ||
            voc_upload(&h->d_w_res, w_res) != cudaSuccess || voc_upload(&h->d_b_res, b_res) != cudaSuccess ||
            voc_upload(&h->d_w_out, w_out) != cudaSuccess || voc_upload(&h->d_b_out, b_out) != cudaSuccess ||
            voc_upload(&h->d_up, up) != cudaSuccess)
            return voc_fail(h, AVC_ERR_CUDA, "uploading the vocoder weights failed: %s", cudaGetErrorString(cudaGetLastError()));
    }
    h->loaded = true;
    return AVC_OK;
}

size_t avc_voc_condition_workspace_bytes(avc_voc_handle h, int U, int T) {
    if (!h || U <= 0 || T <= 0) return 0;
    // two stretched intermediates: (T+4)*5 and (T+4)*25 frames
    return ((size_t)U * V_FEAT * (T + 2 * V_PAD) * 30 * sizeof(float) + 1023) & ~size_t(1023);
}

int avc_voc_condition(avc_voc_handle h, const float* mel, int U, int T, float* m_up, float* aux, void* workspace, void* stream) {
    if (!h || !mel || !m_up || !aux || !workspace) return voc_fail(h, AVC_ERR_INVALID_ARG, "null argument");
    if (h->device < 0) return voc_fail(h, AVC_ERR_CUDA, "host-only handle");
    if (!h->loaded) return voc_fail(h, AVC_ERR_NOT_LOADED, "avc_voc_condition before avc_voc_load_weights");
    if (U <= 0 || T <= 0) return voc_fail(h, AVC_ERR_INVALID_ARG, "U and T must be positive");
    cudaStream_t s = (cudaStream_t)stream;
    ResnetArgs ra{mel, aux, h->d_w_in, h->d_b_in, h->d_w_res, h->d_b_res, h->d_w_out, h->d_b_out, U, T};
    voc_resnet_kernel<<<dim3(ceil_div(T, RF), U), V_CD, 0, s>>>(ra);
    const int Tp = T + 2 * V_PAD;
    float* s1 = (float*)workspace;                               // [U][80][Tp*5]
    float* s2 = s1 + (size_t)U * V_FEAT * Tp * 5;                // [U][80][Tp*25]
    auto grid = [&](long long n) { return (int)std::min<long long>((n + 255) / 256, (long long)h->num_sms * 16); };
    voc_stretch_conv_kernel<<<grid((long long)U * V_FEAT * Tp * 5), 256, 0, s>>>(mel, s1, h->d_up, 5, U, T, V_PAD, 0, (long long)Tp * 5, 0);
    voc_stretch_conv_kernel<<<grid((long long)U * V_FEAT * Tp * 25), 256, 0, s>>>(s1, s2, h->d_up + 11, 5, U, Tp * 5, 0, 0, (long long)Tp * 25, 0);
    voc_stretch_conv_kernel<<<grid((long long)U * V_FEAT * T * V_HOP), 256, 0, s>>>(s2, m_up, h->d_up + 22, 11, U, Tp * 25, 0,
                                                                                   (long long)V_PAD * V_HOP, (long long)T * V_HOP, 1);
    h->launches = 4;
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return voc_fail(h, AVC_ERR_CUDA, "vocoder conditioning launch failed: %s", cudaGetErrorString(e));
    return AVC_OK;
}

int avc_voc_max_rows(void) { return MAX_TILES * TILE; }

size_t avc_voc_generate_workspace_bytes(avc_voc_handle h, int rows, int steps) {
    if (!h || rows <= 0 || steps <= 0) return 0;
    return voc_layout(tiles_for(std::min(rows, MAX_TILES * TILE)), steps).total;
}

int avc_voc_generate(avc_voc_handle h, const float* m_up, const float* aux, int U, int T, const int32_t* row_utt, const int32_t* row_start,
                     int rows, int steps, const float* uniforms, uint64_t seed, const float* forced_x, float* samples, float* logits,
                     void* workspace, void* stream) {
    if (!h || !m_up || !aux || !row_utt || !row_start || !samples || !workspace) return voc_fail(h, AVC_ERR_INVALID_ARG, "null argument");
    if (h->device < 0) return voc_fail(h, AVC_ERR_CUDA, "host-only handle");
    if (!h->loaded) return voc_fail(h, AVC_ERR_NOT_LOADED, "avc_voc_generate before avc_voc_load_weights");
    if (U <= 0 || T <= 0 || steps <= 0 || rows <= 0) return voc_fail(h, AVC_ERR_INVALID_ARG, "U, T, rows and steps must be positive");
    if (rows > MAX_TILES * TILE) return voc_fail(h, AVC_ERR_UNSUPPORTED, "at most %d rows per call (got %d): split the work list", MAX_TILES * TILE, rows);
    cudaStream_t s = (cudaStream_t)stream;
    const int tiles = tiles_for(rows);
    const Layout lay = voc_layout(tiles, steps);
    uint8_t* ws = (uint8_t*)workspace;
    const long long n = (long long)steps * tiles * TILE * COND_BLKS * 8;
    voc_pack_cond_kernel<<<(int)std::min<long long>((n + 255) / 256, (long long)h->num_sms * 32), 256, 0, s>>>(
        m_up, aux, row_utt, row_start, ws + lay.cond, rows, tiles, steps, (long long)T * V_HOP, T);
    cudaMemsetAsync(ws + lay.h1, 0, lay.total - lay.h1, s);      // zero initial state, zero counter
    VocArgs a{};
    a.R = rows; a.L = steps;
    a.wpack = h->d_wpack; a.params = h->d_params; a.cond = ws + lay.cond;
    a.h1 = ws + lay.h1; a.h2 = ws + lay.h2; a.x1 = ws + lay.x1; a.x2 = ws + lay.x2; a.y1 = ws + lay.y1; a.y2 = ws + lay.y2;
    a.counter = (unsigned int*)(ws + lay.counter);
    a.uniforms = uniforms; a.seed = seed; a.forced = forced_x; a.samples = samples; a.logits = logits;
    a.idesc32 = ptx::umma_idesc_f16(1, TILE, 32);
    a.idesc16 = ptx::umma_idesc_f16(1, TILE, 16);
    void* kargs[] = {&a};
    const void* fn = tiles == 1 ? (const void*)voc_generate_kernel<1> : tiles == 2 ? (const void*)voc_generate_kernel<2> : (const void*)voc_generate_kernel<4>;
    a.cs = h->cs;
    a.stages = getenv("AVC_VOC_STAGES") ? std::max(2, std::min(STAGES, atoi(getenv("AVC_VOC_STAGES")))) : STAGES;
    // CTAs wait on one another's output: cooperative launch (guaranteed co-residency), with the checked non-cooperative retry
    // of launch_persistent() for profilers that cannot replay cooperative cluster launches (AVC_VOC_NOCOOP forces it)
    cudaError_t e = launch_persistent(fn, dim3(VCTAS), dim3(256), SMEM_BYTES, (unsigned)h->cs, kargs, s, getenv("AVC_VOC_NOCOOP") != nullptr);
    h->launches = 3;
    if (e != cudaSuccess) return voc_fail(h, AVC_ERR_CUDA, "vocoder loop launch failed: %s", cudaGetErrorString(e));
    return AVC_OK;
}

int64_t avc_voc_last_launch_count(avc_voc_handle h) { return h ? h->launches : 0; }

const char* avc_voc_last_error(avc_voc_handle h) { return h ? h->err.c_str() : g_voc_err.c_str(); }

void avc_voc_destroy(avc_voc_handle h) {
    if (!h) return;
    if (h->device >= 0) {
        cudaSetDevice(h->device);
        for (void* p : {(void*)h->d_wpack, (void*)h->d_params, (void*)h->d_w_in, (void*)h->d_b_in, (void*)h->d_w_res, (void*)h->d_b_res,
                        (void*)h->d_w_out, (void*)h->d_b_out, (void*)h->d_up})
            if (p) cudaFree(p);
    }
    delete h;
}

}  // extern "C"
