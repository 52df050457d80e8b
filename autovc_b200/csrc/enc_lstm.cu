// Encoder BiLSTM recurrence (reference encoder.py:51,74-77: nn.LSTM(512, 32, 2, bidirectional)), H = 32.
//
// The recurrent matrix is tiny (128 x 32 per direction), the chains (utterance x direction) are independent, so
// there is nothing to exchange between CTAs: a group of four warps owns 8 utterances of one direction for the whole
// sequence.  W_hh lives in registers as mma.sync B fragments (16 registers per warp), the cell state and h stay in
// registers, and the unit/gate permutation below makes a thread's accumulator values of step t exactly its A-fragment
// values of step t+1: a step is 8 HMMA + the gate math of two units per thread and one 32-bit exchange between the warps.
//
//   D tile j (8 columns), thread (g = lane/4, q = lane%4) holds columns 2q,2q+1 for row g (rows 8..15 unused)
//   column c of tile j  <->  unit (c/2)*8 + j/2, gate (j%2)*2 + c%2        => thread q owns units q*8 .. q*8+7,
//   K position (s, kl)  <->  unit (kl%8/2)*8 + 4s + (kl/8)*2 + kl%2           all four gates, row g
//
// Layer 1 also performs the reference's code sampling (encoder.py:94-95) and writes the de-duplicated decoder
// input (model.py:115-124) directly: forward h at t = 31 mod 32, backward h at t = 0 mod 32, in fp32.
#include "common.cuh"
#include "launch.h"

namespace avc {

namespace {

constexpr int EH = 32;   // hidden units per direction this kernel is specialised for

// One MUFU.TANH per gate value (sigmoid(x) = 0.5 tanh(0.5 x) + 0.5): the step is a single warp's dependent chain and the
// transcendental unit is its longest stretch (40 values per thread and step).  abs err ~5e-4, below the bf16 rounding of h;
// measured effect on the codes: cosine unchanged to 5 digits (tests/test_gpu_parity.py).
__device__ __forceinline__ float fast_tanh(float x) { float y; asm("tanh.approx.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
__device__ __forceinline__ float fast_sigmoid(float x) { return fmaf(0.5f, fast_tanh(0.5f * x), 0.5f); }

__device__ __forceinline__ void mma_bf16_16816(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
    asm volatile(
        "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
        : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
        : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

__device__ __forceinline__ uint32_t pack_bf16(float lo, float hi) {
    __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
    return *reinterpret_cast<uint32_t*>(&v);
}

struct EncLstmArgs {
    const float* Gx;            // [B][T][2*128] fp32 (direction-major, unit-major gates), bias included
    const __nv_bfloat16* Whh;   // [2][128][32] bf16, row = unit*4 + gate
    __nv_bfloat16* Y;           // layer 0: [B][T][64] bf16 output (fwd | bwd)
    float* codes;               // layer 1: [B][32*(nf+nb)] fp32
    __nv_bfloat16* D;           // layer 1: [B][nb][ldd] bf16 decoder input (fwd code | bwd code | c_trg)
    int ldd;
    int B, T, freq;
};

// A group of NW = 4 warps = 8 utterances (rows g = lane/4 of the m16 tile; rows 8..15 stay zero: the chains are
// latency-bound, so halving the per-thread gate math is worth more than filling the tile) of one direction.  Warp v owns
// units q*8 + 2v, q*8 + 2v + 1 of every thread position: 8 of the 32 HMMAs and 10 of the 40 MUFU ops of a step.  Its packed
// (h[2v], h[2v+1]) is exactly one of the four A-fragment registers of the next step, so the exchange between the warps is one
// 32-bit shared-memory word per thread, one barrier, four 32-bit reads (double-buffered by step parity).  The arithmetic per
// unit is unchanged, so the results are bitwise those of the one-warp-per-chain version (0.52 us/step).
constexpr int NW = 4;
template <bool kCodes>
__global__ void __launch_bounds__(32 * NW) enc_lstm_kernel(const EncLstmArgs a) {
    const int lane = threadIdx.x & 31, wv = threadIdx.x >> 5;
    const int g = lane >> 2, q = lane & 3;
    const int dir = blockIdx.y;
    const int b = blockIdx.x * 8 + g;
    const bool ok = b < a.B;

    // resident B fragments of this warp's tiles j = 4v .. 4v+3 (units 2v, 2v+1: tiles (i,f) and (g,o) each), k-step s
    uint32_t wb[4][2][2];
    const __nv_bfloat16* W = a.Whh + (size_t)dir * 4 * EH * EH;
#pragma unroll
    for (int jj = 0; jj < 4; ++jj) {
        const int j = 4 * wv + jj;
        const int unit = (g >> 1) * 8 + (j >> 1), gate = (j & 1) * 2 + (g & 1);
        const __nv_bfloat16* wr = W + (size_t)(unit * 4 + gate) * EH + q * 8;
#pragma unroll
        for (int s = 0; s < 2; ++s) {
            const uint2 v = *reinterpret_cast<const uint2*>(wr + 4 * s);
            wb[jj][s][0] = v.x;
            wb[jj][s][1] = v.y;
        }
    }

    float c[2] = {0.f, 0.f};
    uint32_t af[2][4];                       // A fragments of h_{t-1} for the two k-steps (a1, a3 = rows 8..15 = 0)
#pragma unroll
    for (int s = 0; s < 2; ++s)
#pragma unroll
        for (int i = 0; i < 4; ++i) af[s][i] = 0u;

    const int nf = a.T / a.freq, nb = (a.T + a.freq - 1) / a.freq;
    const size_t gx_pitch = (size_t)a.T * 8 * EH;            // floats per utterance
    const float* gx = a.Gx + (size_t)(ok ? b : 0) * gx_pitch + dir * 4 * EH + q * 32 + 8 * wv;   // units 2v, 2v+1: 8 floats

    // software pipeline: the hoisted input projections of the next NST-1 steps are in flight (cp.async into a private
    // shared-memory ring, [stage][piece][thread] so that the 16-byte reads are conflict-free) while a step is computed: an L2
    // hit costs 1.1-1.3k cycles on this part, several steps of this kernel
    constexpr int NST = 8;
    __shared__ float4 ring[NST][2][32 * NW];
    __shared__ uint32_t hx[2][NW][32];
    auto prefetch = [&](int step) {
        if (step < a.T) {
            const int tt = dir == 0 ? step : a.T - 1 - step;
            const float* p = gx + (size_t)tt * 8 * EH;
#pragma unroll
            for (int w = 0; w < 2; ++w)
                asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((uint32_t)__cvta_generic_to_shared(&ring[step % NST][w][threadIdx.x])), "l"(p + 4 * w) : "memory");
        }
        asm volatile("cp.async.commit_group;" ::: "memory");      // one group per step, empty past the end
    };
#pragma unroll
    for (int i = 0; i < NST - 1; ++i) prefetch(i);

    // t % freq and t / freq of the running frame, kept incrementally (a runtime division per step sits in the single
    // instruction stream of the chain)
    int tm = dir == 0 ? 0 : (a.T - 1) % a.freq, tj = dir == 0 ? 0 : (a.T - 1) / a.freq;
    for (int step = 0; step < a.T; ++step) {
        prefetch(step + NST - 1);
        asm volatile("cp.async.wait_group %0;" ::"n"(NST - 1) : "memory");   // this thread's pieces of `step` have landed
        const int t = dir == 0 ? step : a.T - 1 - step;
        float h[2];
#pragma unroll
        for (int w = 0; w < 2; ++w) {
            const float4 x = ring[step % NST][w][threadIdx.x];
            // tiles (i,f) and (g,o) of unit q*8 + 2v + w
            float d_if[4] = {0.f, 0.f, 0.f, 0.f}, d_go[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
            for (int s = 0; s < 2; ++s) {
                mma_bf16_16816(d_if, af[s], wb[2 * w][s][0], wb[2 * w][s][1]);
                mma_bf16_16816(d_go, af[s], wb[2 * w + 1][s][0], wb[2 * w + 1][s][1]);
            }
            const float ig = fast_sigmoid(d_if[0] + x.x), fg = fast_sigmoid(d_if[1] + x.y);
            const float gg = fast_tanh(d_go[0] + x.z), og = fast_sigmoid(d_go[1] + x.w);
            c[w] = fg * c[w] + ig * gg;
            h[w] = og * fast_tanh(c[w]);
        }
        // exchange: warp v's pair is A-fragment register (k-step v/2, a0 or a2) of everybody's next step
        hx[step & 1][wv][lane] = pack_bf16(h[0], h[1]);
        __syncthreads();
        af[0][0] = hx[step & 1][0][lane];
        af[0][2] = hx[step & 1][1][lane];
        af[1][0] = hx[step & 1][2][lane];
        af[1][2] = hx[step & 1][3][lane];
        const uint4 v = make_uint4(af[0][0], af[0][2], af[1][0], af[1][2]);
        if constexpr (!kCodes) {
            // layer 0: h_t (bf16) -> Y[b][t][dir*32 + q*8 .. +7]; the warps take turns
            if (ok && wv == (step & (NW - 1))) *reinterpret_cast<uint4*>(a.Y + ((size_t)b * a.T + t) * (2 * EH) + dir * EH + q * 8) = v;
        } else {
            // layer 1: code sampling (fp32, every warp its own two units) + de-duplicated decoder input (bf16, warp 0)
            const bool hit = dir == 0 ? (tm == a.freq - 1) : (tm == 0);
            if (hit && ok) {
                const int j = tj;
                const size_t cl = (size_t)EH * (nf + nb);
                const size_t coff = (dir == 0 ? (size_t)j * EH : (size_t)(nf + j) * EH) + q * 8 + 2 * wv;
                *reinterpret_cast<float2*>(a.codes + (size_t)b * cl + coff) = make_float2(h[0], h[1]);
                if (wv == 0) {
                    __nv_bfloat16* dp = a.D + ((size_t)b * nb + j) * a.ldd + dir * EH + q * 8;
                    *reinterpret_cast<uint4*>(dp) = v;
                    // the tail segment (t >= freq*nf) repeats the LAST forward code (reference model.py:116-117)
                    if (dir == 0 && j == nf - 1 && nb > nf) *reinterpret_cast<uint4*>(dp + a.ldd) = v;
                }
            }
            if (dir == 0) { if (++tm == a.freq) { tm = 0; ++tj; } }
            else { if (tm == 0) { tm = a.freq - 1; --tj; } else --tm; }
        }
    }
}

// c_trg (B,E) fp32 -> D[b][s][2*dn .. 2*dn+E) bf16 for every segment s (speaker part of the decoder input)
__global__ void fill_speaker_kernel(const float* __restrict__ c_trg, __nv_bfloat16* __restrict__ D, int nb, int ldd, int off, int E) {
    const int b = blockIdx.x;
    for (int i = threadIdx.x; i < nb * E; i += blockDim.x) {
        const int s = i / E, e = i - s * E;
        D[((size_t)b * nb + s) * ldd + off + e] = __float2bfloat16_rn(c_trg[(size_t)b * E + e]);
    }
}

}  // namespace

bool enc_lstm_supported(int dim_neck) { return dim_neck == EH; }

cudaError_t launch_enc_lstm_l0(const float* Gx, const void* Whh_bf16, void* Y_bf16, int B, int T, cudaStream_t s) {
    EncLstmArgs a{};
    a.Gx = Gx; a.Whh = (const __nv_bfloat16*)Whh_bf16; a.Y = (__nv_bfloat16*)Y_bf16; a.B = B; a.T = T; a.freq = 1;
    enc_lstm_kernel<false><<<dim3(ceil_div(B, 8), 2), 32 * NW, 0, s>>>(a);
    return cudaGetLastError();
}

cudaError_t launch_enc_lstm_l1(const float* Gx, const void* Whh_bf16, float* codes, void* D_bf16, int ldd, const float* c_trg,
                               int E, int B, int T, int freq, cudaStream_t s, int* launches) {
    EncLstmArgs a{};
    a.Gx = Gx; a.Whh = (const __nv_bfloat16*)Whh_bf16; a.codes = codes; a.D = (__nv_bfloat16*)D_bf16; a.ldd = ldd;
    a.B = B; a.T = T; a.freq = freq;
    const int nb = ceil_div(T, freq);
    if (!D_bf16) return cudaErrorInvalidValue;   // the caller always provides the workspace segment matrix
    if (c_trg) {                                  // encode-only passes (c_trg == nullptr) skip the speaker part
        fill_speaker_kernel<<<B, 256, 0, s>>>(c_trg, a.D, nb, ldd, 2 * EH, E);
        if (launches) ++*launches;
    }
    enc_lstm_kernel<true><<<dim3(ceil_div(B, 8), 2), 32 * NW, 0, s>>>(a);
    if (launches) ++*launches;
    return cudaGetLastError();
}

}  // namespace avc
