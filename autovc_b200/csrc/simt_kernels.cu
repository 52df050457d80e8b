// CUDA-core (SIMT) kernels: the fp32 validation path of every layer, plus the small layout kernels
// (input packing, code sampling) shared by both precisions.
//
// The validation mode exists to hold the reference to <= 1e-4 (north-star) - fp32 operands, fp32 FFMA
// accumulation, accurate expf/tanhf - and as the on-device cross-check for the tcgen05 path.
#include "common.cuh"
#include "launch.h"

namespace avc {

// ---------------------------------------------------------------------------------------------
// element helpers
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ float4 load4(const float* p) { return *reinterpret_cast<const float4*>(p); }
__device__ __forceinline__ float4 load4(const __nv_bfloat16* p) {
    uint2 r = *reinterpret_cast<const uint2*>(p);
    __nv_bfloat162 a = *reinterpret_cast<__nv_bfloat162*>(&r.x), b = *reinterpret_cast<__nv_bfloat162*>(&r.y);
    float2 fa = __bfloat1622float2(a), fb = __bfloat1622float2(b);
    return make_float4(fa.x, fa.y, fb.x, fb.y);
}
__device__ __forceinline__ float4 load4(const __half* p) {
    uint2 r = *reinterpret_cast<const uint2*>(p);
    __half2 a = *reinterpret_cast<__half2*>(&r.x), b = *reinterpret_cast<__half2*>(&r.y);
    float2 fa = __half22float2(a), fb = __half22float2(b);
    return make_float4(fa.x, fa.y, fb.x, fb.y);
}
__device__ __forceinline__ void store4(float* p, float4 v) { *reinterpret_cast<float4*>(p) = v; }
__device__ __forceinline__ void store4(__nv_bfloat16* p, float4 v) {
    __nv_bfloat162 a = __floats2bfloat162_rn(v.x, v.y), b = __floats2bfloat162_rn(v.z, v.w);
    uint2 r;
    r.x = *reinterpret_cast<uint32_t*>(&a);
    r.y = *reinterpret_cast<uint32_t*>(&b);
    *reinterpret_cast<uint2*>(p) = r;
}
__device__ __forceinline__ void store4(__half* p, float4 v) {
    __half2 a = __floats2half2_rn(v.x, v.y), b = __floats2half2_rn(v.z, v.w);
    uint2 r;
    r.x = *reinterpret_cast<uint32_t*>(&a);
    r.y = *reinterpret_cast<uint32_t*>(&b);
    *reinterpret_cast<uint2*>(p) = r;
}
__device__ __forceinline__ void store1(float* p, float v) { *p = v; }
__device__ __forceinline__ void store1(__nv_bfloat16* p, float v) { *p = __float2bfloat16_rn(v); }
__device__ __forceinline__ void store1(__half* p, float v) { *p = __float2half_rn(v); }

__device__ __forceinline__ float sigmoid_acc(float x) { return 1.0f / (1.0f + expf(-x)); }

__device__ __forceinline__ float apply_act(float v, int act) {
    if (act == ACT_RELU) return v < 0.0f ? 0.0f : v;          // like torch.relu: a NaN stays a NaN (fmaxf would return 0)
    if (act == ACT_TANH) return tanhf(v);
    return v;
}

// ---------------------------------------------------------------------------------------------
// 64x64x16 register-tiled FFMA main loop (256 threads, 4x4 per thread, double-buffered smem)
// ---------------------------------------------------------------------------------------------
constexpr int BM = 64, BN = 64, BK = 16, PITCH = 68;

template <typename TW, typename ALoad>
__device__ __forceinline__ void simt_mainloop(float (&acc)[4][4], ALoad a_load, const TW* __restrict__ W, int ldw,
                                              int n0, int N, int K) {
    __shared__ __align__(16) float As[2][BK][PITCH];
    __shared__ __align__(16) float Ws[2][BK][PITCH];
    const int tid = threadIdx.x;
    const int lr = tid >> 2;         // tile row (A) / tile column (W) this thread stages
    const int kq = (tid & 3) * 4;    // k offset inside the BK slab
    const int tx = tid & 15, ty = tid >> 4;
    const bool w_ok = (n0 + lr) < N;
    const TW* wrow = W + (size_t)(n0 + lr) * ldw;

#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = 0.0f;

    float4 ra = a_load(kq);
    float4 rw = w_ok ? load4(wrow + kq) : make_float4(0.f, 0.f, 0.f, 0.f);
    int buf = 0;
    As[0][kq + 0][lr] = ra.x; As[0][kq + 1][lr] = ra.y; As[0][kq + 2][lr] = ra.z; As[0][kq + 3][lr] = ra.w;
    Ws[0][kq + 0][lr] = rw.x; Ws[0][kq + 1][lr] = rw.y; Ws[0][kq + 2][lr] = rw.z; Ws[0][kq + 3][lr] = rw.w;
    __syncthreads();

    for (int k0 = 0; k0 < K; k0 += BK) {
        const bool more = (k0 + BK) < K;
        if (more) {
            ra = a_load(k0 + BK + kq);
            rw = w_ok ? load4(wrow + k0 + BK + kq) : make_float4(0.f, 0.f, 0.f, 0.f);
        }
#pragma unroll
        for (int k = 0; k < BK; ++k) {
            const float4 a = *reinterpret_cast<const float4*>(&As[buf][k][ty * 4]);
            const float4 w = *reinterpret_cast<const float4*>(&Ws[buf][k][tx * 4]);
            const float av[4] = {a.x, a.y, a.z, a.w};
            const float wv[4] = {w.x, w.y, w.z, w.w};
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(av[i], wv[j], acc[i][j]);
        }
        if (more) {
            const int nb = buf ^ 1;
            As[nb][kq + 0][lr] = ra.x; As[nb][kq + 1][lr] = ra.y; As[nb][kq + 2][lr] = ra.z; As[nb][kq + 3][lr] = ra.w;
            Ws[nb][kq + 0][lr] = rw.x; Ws[nb][kq + 1][lr] = rw.y; Ws[nb][kq + 2][lr] = rw.z; Ws[nb][kq + 3][lr] = rw.w;
            __syncthreads();
            buf = nb;
        }
    }
}

// ---------------------------------------------------------------------------------------------
// GEMM / implicit conv1d with fused bias + activation (+ channel-first fp32 output + residual)
// ---------------------------------------------------------------------------------------------
template <typename TA, typename TW, typename TO>
__global__ void __launch_bounds__(256) simt_gemm_kernel(GemmParams p) {
    const int m0 = blockIdx.x * BM, n0 = blockIdx.y * BN;
    const int tid = threadIdx.x;
    const int lr = tid >> 2;
    const int m = m0 + lr;
    const int t_row = (p.taps > 1) ? (m % p.T) : 0;
    const int pad = p.taps / 2;
    const TA* A = reinterpret_cast<const TA*>(p.A);
    const bool m_ok = m < p.M;

    auto a_load = [&](int k) -> float4 {
        if (!m_ok) return make_float4(0.f, 0.f, 0.f, 0.f);
        int tap = 0, ci = k;
        if (p.taps > 1) { tap = k / p.Cp; ci = k - tap * p.Cp; }
        if (ci >= p.Cin) return make_float4(0.f, 0.f, 0.f, 0.f);
        const int tt = t_row + tap - pad;
        if (p.taps > 1 && (tt < 0 || tt >= p.T)) return make_float4(0.f, 0.f, 0.f, 0.f);
        return load4(A + (size_t)(m + tap - pad) * p.lda + ci);
    };

    float acc[4][4];
    simt_mainloop<TW>(acc, a_load, reinterpret_cast<const TW*>(p.W), p.ldw, n0, p.N, p.K);

    const int tx = tid & 15, ty = tid >> 4;
    const int n = n0 + tx * 4;
    if (n >= p.N) return;
    const float4 bv = p.bias ? *reinterpret_cast<const float4*>(p.bias + n) : make_float4(0.f, 0.f, 0.f, 0.f);
    const float bb[4] = {bv.x, bv.y, bv.z, bv.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int mm = m0 + ty * 4 + i;
        if (mm >= p.M) continue;
        float v[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) v[j] = apply_act(acc[i][j] + bb[j], p.act);
        if (p.out) store4(reinterpret_cast<TO*>(p.out) + (size_t)mm * p.ldo + n, make_float4(v[0], v[1], v[2], v[3]));
        if (p.out_t) {
            const int b = mm / p.T, t = mm - b * p.T;
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const size_t o = ((size_t)b * p.N + (n + j)) * p.T + t;
                p.out_t[o] = v[j] + (p.res_t ? p.res_t[o] : 0.0f);
            }
        }
    }
}

// ---------------------------------------------------------------------------------------------
// One LSTM time step (recurrent GEMM + fused gate math), both directions via blockIdx.z
// ---------------------------------------------------------------------------------------------
template <typename TH, typename TW>
__global__ void __launch_bounds__(256) simt_lstm_step_kernel(LstmStepParams q) {
    const int d = blockIdx.z;
    const int m0 = blockIdx.x * BM, n0 = blockIdx.y * BN;
    const int tid = threadIdx.x;
    const int lr = tid >> 2;
    const int b_ld = m0 + lr;
    const TH* hp = reinterpret_cast<const TH*>(q.h_prev[d]);
    const bool ok = (b_ld < q.B) && (hp != nullptr);

    auto a_load = [&](int k) -> float4 {
        if (!ok) return make_float4(0.f, 0.f, 0.f, 0.f);
        return load4(hp + (size_t)b_ld * q.ldh + k);
    };
    float acc[4][4];
    if (hp != nullptr) {
        simt_mainloop<TW>(acc, a_load, reinterpret_cast<const TW*>(q.Whh[d]), q.H, n0, 4 * q.H, q.H);
    } else {
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
            for (int j = 0; j < 4; ++j) acc[i][j] = 0.0f;
    }

    const int tx = tid & 15, ty = tid >> 4;
    const int n = n0 + tx * 4;
    const int u = n >> 2;
    if (u >= q.H) return;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int b = m0 + ty * 4 + i;
        if (b >= q.B) continue;
        const float4 gx = *reinterpret_cast<const float4*>(q.Gx[d] + (size_t)b * q.ldg + n);
        const float ig = sigmoid_acc(acc[i][0] + gx.x);
        const float fg = sigmoid_acc(acc[i][1] + gx.y);
        const float gg = tanhf(acc[i][2] + gx.z);
        const float og = sigmoid_acc(acc[i][3] + gx.w);
        float* cp = q.c[d] + (size_t)b * q.H + u;
        const float c_new = fg * (*cp) + ig * gg;
        *cp = c_new;
        store1(reinterpret_cast<TH*>(q.h_out[d]) + (size_t)b * q.ldho + u, og * tanhf(c_new));
    }
}

// ---------------------------------------------------------------------------------------------
// Layout kernels
// ---------------------------------------------------------------------------------------------
// x (B,80,T) fp32 channel-first + c_org (B,E) -> X0 (B,T,80+E) channel-last of type TO.
// reference encoder.py:65-66 (expand + cat) without materialising the fp32 concat.
template <typename TO>
__global__ void pack_input_kernel(const float* __restrict__ x, const float* __restrict__ c_org, TO* __restrict__ X0,
                                  int B, int T, int E) {
    __shared__ float tile[32][33];
    const int b = blockIdx.z;
    const int t0 = blockIdx.x * 32;
    const int C = N_MELS + E;
    const int tx = threadIdx.x, ty = threadIdx.y;  // 32 x 8
    // mel part: transpose 32(t) x 32(c) tiles
    for (int c0 = 0; c0 < N_MELS; c0 += 32) {
        for (int j = ty; j < 32; j += 8) {
            const int c = c0 + j, t = t0 + tx;
            tile[j][tx] = (c < N_MELS && t < T) ? x[((size_t)b * N_MELS + c) * T + t] : 0.0f;
        }
        __syncthreads();
        for (int j = ty; j < 32; j += 8) {
            const int t = t0 + j, c = c0 + tx;
            if (t < T && c < N_MELS) store1(X0 + ((size_t)b * T + t) * C + c, tile[tx][j]);
        }
        __syncthreads();
    }
    // speaker part: broadcast over t
    for (int j = ty; j < 32; j += 8) {
        const int t = t0 + j;
        if (t >= T) continue;
        for (int e = tx; e < E; e += 32) store1(X0 + ((size_t)b * T + t) * C + N_MELS + e, c_org[(size_t)b * E + e]);
    }
}

// Encoder output Y (B,T,2*dn) fp32 -> content codes (reference encoder.py:94-95, model.py:99,143) and,
// when D != nullptr, the de-duplicated decoder input D (B, nb, 2*dn+E): one row per distinct column of the
// reference's up-sampled (B,320,T) tensor (model.py:115-124) - fwd code min(s,nf-1) | bwd code s | c_trg.
template <typename TO>
__global__ void codes_kernel(const float* __restrict__ Y, const float* __restrict__ c_trg, float* __restrict__ codes,
                             TO* __restrict__ D, int B, int T, int dn, int E, int freq) {
    const int b = blockIdx.x;
    const int nf = T / freq, nb = (T + freq - 1) / freq;
    const int ldy = 2 * dn;
    for (int i = threadIdx.x; i < nf * dn; i += blockDim.x) {
        const int j = i / dn, u = i - j * dn;
        codes[(size_t)b * dn * (nf + nb) + i] = Y[((size_t)b * T + (j * freq + freq - 1)) * ldy + u];
    }
    for (int i = threadIdx.x; i < nb * dn; i += blockDim.x) {
        const int j = i / dn, u = i - j * dn;
        codes[(size_t)b * dn * (nf + nb) + nf * dn + i] = Y[((size_t)b * T + j * freq) * ldy + dn + u];
    }
    if (D == nullptr) return;
    const int CD = 2 * dn + E;
    for (int i = threadIdx.x; i < nb * CD; i += blockDim.x) {
        const int s = i / CD, c = i - s * CD;
        float v;
        if (c < dn) {
            const int j = min(s, nf - 1);
            v = Y[((size_t)b * T + (j * freq + freq - 1)) * ldy + c];
        } else if (c < 2 * dn) {
            v = Y[((size_t)b * T + s * freq) * ldy + c];
        } else {
            v = c_trg[(size_t)b * E + (c - 2 * dn)];
        }
        store1(D + ((size_t)b * nb + s) * CD + c, v);
    }
}

// ---------------------------------------------------------------------------------------------
// host launchers
// ---------------------------------------------------------------------------------------------
template <typename TA, typename TW>
static cudaError_t launch_simt_gemm_o(const GemmParams& p, cudaStream_t s) {
    dim3 grid(ceil_div(p.M, BM), ceil_div(p.N, BN));
    switch (p.out_type) {
        case DT_F32: simt_gemm_kernel<TA, TW, float><<<grid, 256, 0, s>>>(p); break;
        case DT_BF16: simt_gemm_kernel<TA, TW, __nv_bfloat16><<<grid, 256, 0, s>>>(p); break;
        default: simt_gemm_kernel<TA, TW, __half><<<grid, 256, 0, s>>>(p); break;
    }
    return cudaGetLastError();
}

cudaError_t launch_simt_gemm(const GemmParams& p, cudaStream_t s) {
    if (p.a_type == DT_F32 && p.w_type == DT_F32) return launch_simt_gemm_o<float, float>(p, s);
    if (p.a_type == DT_BF16 && p.w_type == DT_BF16) return launch_simt_gemm_o<__nv_bfloat16, __nv_bfloat16>(p, s);
    if (p.a_type == DT_F16 && p.w_type == DT_F16) return launch_simt_gemm_o<__half, __half>(p, s);
    if (p.a_type == DT_F32 && p.w_type == DT_BF16) return launch_simt_gemm_o<float, __nv_bfloat16>(p, s);
    return cudaErrorInvalidValue;
}

cudaError_t launch_simt_lstm_step(const LstmStepParams& q, int ndir, cudaStream_t s) {
    dim3 grid(ceil_div(q.B, BM), ceil_div(4 * q.H, BN), ndir);
    if (q.h_type == DT_F32 && q.w_type == DT_F32) simt_lstm_step_kernel<float, float><<<grid, 256, 0, s>>>(q);
    else if (q.h_type == DT_BF16 && q.w_type == DT_BF16) simt_lstm_step_kernel<__nv_bfloat16, __nv_bfloat16><<<grid, 256, 0, s>>>(q);
    else if (q.h_type == DT_F32 && q.w_type == DT_BF16) simt_lstm_step_kernel<float, __nv_bfloat16><<<grid, 256, 0, s>>>(q);
    else return cudaErrorInvalidValue;
    return cudaGetLastError();
}

// Encoder conv0's speaker term (reference encoder.py:65-70: the embedding is concatenated on every frame, so its share of the
// 5-tap convolution is constant in t except where the window reaches into the zero padding).  S[b][tap][n] = sum_c W0'[n, 80+c,
// tap] c_org[b, c] comes from one small GEMM; this kernel turns it into the five per-utterance bias vectors the conv0 epilogue adds:
//   E[b][0] interior = bias + S0+S1+S2+S3+S4,   E[b][1] (t = 0) = bias + S2+S3+S4,     E[b][2] (t = 1)   = bias + S1+S2+S3+S4,
//   E[b][3] (t = T-2) = bias + S0+S1+S2+S3,     E[b][4] (t = T-1) = bias + S0+S1+S2.
__global__ void speaker_edge_kernel(const float* __restrict__ S, const float* __restrict__ bias, float* __restrict__ E, int B, int N) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= B * N) return;
    const int b = i / N, n = i - b * N;
    const float* s = S + (size_t)b * 5 * N + n;
    const float s0 = s[0], s1 = s[N], s2 = s[2 * N], s3 = s[3 * N], s4 = s[4 * N], bb = bias[n];
    float* e = E + (size_t)b * 5 * N + n;
    const float mid = s1 + s2 + s3;
    e[0] = bb + (s0 + mid + s4);
    e[N] = bb + (s2 + s3 + s4);
    e[2 * N] = bb + (mid + s4);
    e[3 * N] = bb + (s0 + mid);
    e[4 * N] = bb + (s0 + s1 + s2);
}

// Encoder input for the fused-concat path: the 80 mel channels of x (B,80,T) fp32 channel-first -> X0 [B][T][80] bf16 channel-last
// (one 16-byte store of 8 channels per thread; the loads of a warp run along t), and - same launch, the blocks behind the mel part -
// c_org [B][E] fp32 -> bf16 for the speaker-term GEMM.  Replaces pack_input_kernel (32x32 shared-memory transposes, 11 us for 5 MB)
// + cast_pad_rows on that path.  reference encoder.py:65-66 (what is left of the concat)
__global__ void __launch_bounds__(256) pack_mel_kernel(const float* __restrict__ x, const float* __restrict__ c_org, __nv_bfloat16* __restrict__ X0,
                                                       __nv_bfloat16* __restrict__ Cb, int B, int T, int E, int mel_blocks) {
    if ((int)blockIdx.x < mel_blocks) {
        const long long i = (long long)blockIdx.x * 256 + threadIdx.x;          // (b, channel group of 8, t), t fastest
        const long long per_b = (long long)(N_MELS / 8) * T;
        if (i >= per_b * B) return;
        const int b = (int)(i / per_b), r = (int)(i - (long long)b * per_b), cg = r / T, t = r - cg * T;
        const float* src = x + ((size_t)b * N_MELS + cg * 8) * T + t;
        uint32_t pk[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            __nv_bfloat162 h2 = __floats2bfloat162_rn(src[(size_t)(2 * j) * T], src[(size_t)(2 * j + 1) * T]);
            pk[j] = *reinterpret_cast<uint32_t*>(&h2);
        }
        *reinterpret_cast<uint4*>(X0 + ((size_t)b * T + t) * N_MELS + cg * 8) = make_uint4(pk[0], pk[1], pk[2], pk[3]);
    } else {
        const long long i = ((long long)(blockIdx.x - mel_blocks) * 256 + threadIdx.x) * 2;
        if (i >= (long long)B * E) return;
        *reinterpret_cast<__nv_bfloat162*>(Cb + i) = __floats2bfloat162_rn(c_org[i], c_org[i + 1]);
    }
}

cudaError_t launch_pack_input(const float* x, const float* c_org, void* X0, int out_type, int B, int T, int E,
                              cudaStream_t s) {
    dim3 grid(ceil_div(T, 32), 1, B), block(32, 8);
    if (out_type == DT_F32) pack_input_kernel<float><<<grid, block, 0, s>>>(x, c_org, (float*)X0, B, T, E);
    else if (out_type == DT_BF16) pack_input_kernel<__nv_bfloat16><<<grid, block, 0, s>>>(x, c_org, (__nv_bfloat16*)X0, B, T, E);
    else pack_input_kernel<__half><<<grid, block, 0, s>>>(x, c_org, (__half*)X0, B, T, E);
    return cudaGetLastError();
}

cudaError_t launch_pack_mel(const float* x, const float* c_org, void* X0_bf16, void* Cb_bf16, int B, int T, int E, cudaStream_t s) {
    const long long n_mel = (long long)B * (N_MELS / 8) * T, n_c = ((long long)B * E + 1) / 2;
    const int mel_blocks = (int)((n_mel + 255) / 256), c_blocks = (int)((n_c + 255) / 256);
    pack_mel_kernel<<<mel_blocks + c_blocks, 256, 0, s>>>(x, c_org, (__nv_bfloat16*)X0_bf16, (__nv_bfloat16*)Cb_bf16, B, T, E, mel_blocks);
    return cudaGetLastError();
}

cudaError_t launch_speaker_edge(const float* S, const float* bias, float* E, int B, int N, cudaStream_t s) {
    speaker_edge_kernel<<<ceil_div(B * N, 256), 256, 0, s>>>(S, bias, E, B, N);
    return cudaGetLastError();
}

cudaError_t launch_codes(const float* Y, const float* c_trg, float* codes, void* D, int d_type, int B, int T, int dn,
                         int E, int freq, cudaStream_t s) {
    if (d_type == DT_F32) codes_kernel<float><<<B, 256, 0, s>>>(Y, c_trg, codes, (float*)D, B, T, dn, E, freq);
    else if (d_type == DT_BF16) codes_kernel<__nv_bfloat16><<<B, 256, 0, s>>>(Y, c_trg, codes, (__nv_bfloat16*)D, B, T, dn, E, freq);
    else codes_kernel<__half><<<B, 256, 0, s>>>(Y, c_trg, codes, (__half*)D, B, T, dn, E, freq);
    return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------
// d-vector speaker encoder helpers (reference autovc/speaker_encoder/model.py:70-90, 185-186)
// ---------------------------------------------------------------------------------------------
template <typename TO>
__global__ void cast_pad_rows_kernel(const float* __restrict__ in, TO* __restrict__ out, long long rows, int K, int Kp) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= rows * Kp) return;
    const long long r = i / Kp; const int k = (int)(i - r * Kp);
    const float v = k < K ? in[r * K + k] : 0.0f;
    if constexpr (sizeof(TO) == 4) out[i] = v;
    else out[i] = TO(v);
}

// out[r, :] = in[r, :] / ||in[r, :]||_2 ; one warp per row
__global__ void l2norm_rows_kernel(const float* __restrict__ in, float* __restrict__ out, int rows, int D) {
    const int r = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (r >= rows) return;
    float ss = 0.0f;
    for (int d = lane; d < D; d += 32) { const float v = in[(size_t)r * D + d]; ss += v * v; }
#pragma unroll
    for (int o = 16; o; o >>= 1) ss += __shfl_xor_sync(0xffffffffu, ss, o);
    const float inv = 1.0f / sqrtf(ss);
    for (int d = lane; d < D; d += 32) out[(size_t)r * D + d] = in[(size_t)r * D + d] * inv;
}

// speaker s: mean of its partial embeddings rows [off[s], off[s+1]), re-normalised (embed_utterance, model.py:185-186)
__global__ void segment_mean_renorm_kernel(const float* __restrict__ emb, const int* __restrict__ off, float* __restrict__ out, int D) {
    extern __shared__ float sh[];
    const int s = blockIdx.x, lo = off[s], hi = off[s + 1];
    float ss = 0.0f;
    for (int d = threadIdx.x; d < D; d += blockDim.x) {
        float a = 0.0f;
        for (int r = lo; r < hi; ++r) a += emb[(size_t)r * D + d];
        a /= (float)(hi - lo);
        sh[d] = a;
        ss += a * a;
    }
    __shared__ float red[32];
#pragma unroll
    for (int o = 16; o; o >>= 1) ss += __shfl_xor_sync(0xffffffffu, ss, o);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = ss;
    __syncthreads();
    float tot = 0.0f;
    for (int w = 0; w < (int)(blockDim.x >> 5); ++w) tot += red[w];
    const float inv = 1.0f / sqrtf(tot);
    for (int d = threadIdx.x; d < D; d += blockDim.x) out[(size_t)s * D + d] = sh[d] * inv;
}

cudaError_t launch_cast_pad_rows(const float* in, void* out, int out_type, long long rows, int K, int Kp, cudaStream_t s) {
    const long long n = rows * Kp;
    const unsigned grid = (unsigned)((n + 255) / 256);
    if (out_type == DT_F32) cast_pad_rows_kernel<float><<<grid, 256, 0, s>>>(in, (float*)out, rows, K, Kp);
    else if (out_type == DT_BF16) cast_pad_rows_kernel<__nv_bfloat16><<<grid, 256, 0, s>>>(in, (__nv_bfloat16*)out, rows, K, Kp);
    else cast_pad_rows_kernel<__half><<<grid, 256, 0, s>>>(in, (__half*)out, rows, K, Kp);
    return cudaGetLastError();
}

cudaError_t launch_l2norm_rows(const float* in, float* out, int rows, int D, cudaStream_t s) {
    l2norm_rows_kernel<<<ceil_div(rows, 8), 256, 0, s>>>(in, out, rows, D);
    return cudaGetLastError();
}

cudaError_t launch_segment_mean_renorm(const float* emb, const int* offsets, float* out, int S, int D, cudaStream_t s) {
    segment_mean_renorm_kernel<<<S, 256, D * sizeof(float), s>>>(emb, offsets, out, D);
    return cudaGetLastError();
}

}  // namespace avc
