// Train-mode BatchNorm1d forward (batch statistics) for the conv stacks: SURVEY 8f-4.
//
// Reference: nn.BatchNorm1d(512 / 80) in training mode as the un-eval()'d module runs it
// (autovc/auto_encoder/encoder.py:46,70, decoder.py:48,70, postnet.py:27,38,47,55,61; model.py:306,201):
//   mean_c = mean over (B,T) of z[:,c,:],  var_c = biased variance over (B,T),
//   y = (z - mean_c) / sqrt(var_c + 1e-5) * gamma_c + beta_c, then ReLU / tanh / nothing.
// z = conv(x) + bias comes from the GEMM kernels as fp32 rows [M = B*T][N] (channel-last), so the statistics are a
// column reduction.  Three HBM-bound launches per layer:
//   bn_partial   rows are cut into chunks; each block writes per-channel (sum, sum of squares) of its chunk in
//                double precision (fixed order inside the block: bitwise reproducible, no atomics)
//   bn_finalize  sums the chunk partials in chunk order -> mean, biased var, scale = gamma/sqrt(var+eps),
//                shift = beta - mean*scale; also hands (mean, var) back for the running-statistics update
//   bn_apply     y = act(z*scale + shift) -> next layer's operand (bf16 / fp16 / fp32 rows), or the reference's
//                channel-first fp32 (B,80,T) with the postnet residual for the last layer
// Algorithmic bytes per layer and frame (N channels): 4N (z read for the statistics) + 4N (z read to apply) + 2N out.
#include "common.cuh"
#include "launch.h"

namespace avc {

namespace {

constexpr int BN_THREADS = 256;

// partial[chunk][0][c] = sum_r z[r,c], partial[chunk][1][c] = sum_r z[r,c]^2 over the chunk's rows
__global__ void __launch_bounds__(BN_THREADS) bn_partial_kernel(const float* __restrict__ Z, double* __restrict__ partial,
                                                                int M, int N, int rows_per_chunk) {
    extern __shared__ double sh[];                       // [rows_per_pass][2][N]
    const int cols4 = N >> 2;                            // N <= 1024: one float4 column per thread
    const int rpp = BN_THREADS / cols4;                  // rows handled per pass
    const int r0 = blockIdx.x * rows_per_chunk;
    const int r1 = min(M, r0 + rows_per_chunk);
    const int rp = threadIdx.x / cols4, c4 = threadIdx.x - rp * cols4;
    if (rp < rpp) {
        double s[4] = {0, 0, 0, 0}, q[4] = {0, 0, 0, 0};
        auto acc = [&](const float4& v) {
            s[0] += v.x; q[0] += (double)v.x * v.x;
            s[1] += v.y; q[1] += (double)v.y * v.y;
            s[2] += v.z; q[2] += (double)v.z * v.z;
            s[3] += v.w; q[3] += (double)v.w * v.w;
        };
        const float* zp = Z + 4 * c4;
        int r = r0 + rp;
        for (; r + 3 * rpp < r1; r += 4 * rpp) {          // four rows in flight per thread: the pass is latency-bound otherwise
            const float4 v0 = *reinterpret_cast<const float4*>(zp + (size_t)r * N);
            const float4 v1 = *reinterpret_cast<const float4*>(zp + (size_t)(r + rpp) * N);
            const float4 v2 = *reinterpret_cast<const float4*>(zp + (size_t)(r + 2 * rpp) * N);
            const float4 v3 = *reinterpret_cast<const float4*>(zp + (size_t)(r + 3 * rpp) * N);
            acc(v0); acc(v1); acc(v2); acc(v3);
        }
        for (; r < r1; r += rpp) acc(*reinterpret_cast<const float4*>(zp + (size_t)r * N));
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            sh[((size_t)rp * 2 + 0) * N + 4 * c4 + j] = s[j];
            sh[((size_t)rp * 2 + 1) * N + 4 * c4 + j] = q[j];
        }
    }
    __syncthreads();
    for (int i = threadIdx.x; i < 2 * N; i += BN_THREADS) {
        double a = 0;
        for (int p = 0; p < rpp; ++p) a += sh[(size_t)p * 2 * N + i];
        partial[(size_t)blockIdx.x * 2 * N + i] = a;
    }
}

// 32 channels per block, 32 warps split the chunks; the cross-warp sum runs in warp order
// FUSED: the chunks come from the GEMM epilogue (tc_gemm.cu, STATS): four floats per (chunk, channel) - S1 = sum(x - shift),
// S2 = sum((x - shift)^2), the chunk's own shift (its first valid row) and its number of valid rows - which are re-based here,
// in double, onto the shift g of chunk 0:  d = shift - g,  S1' = S1 + n d,  S2' = S2 + 2 d S1 + n d^2.  Otherwise the chunks are
// bn_partial's plain double-precision (sum, sum of squares).
constexpr int FIN_WARPS = 32;
template <bool FUSED>
__global__ void __launch_bounds__(FIN_WARPS * 32) bn_finalize_kernel(const double* __restrict__ partial, int chunks, int M, int N,
                                                                 const float* __restrict__ gamma, const float* __restrict__ beta,
                                                                 float eps, float* __restrict__ scale_shift, float* __restrict__ stat_mean,
                                                                 float* __restrict__ stat_var) {
    __shared__ double sh[FIN_WARPS][2][32];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int c = blockIdx.x * 32 + lane;
    double s = 0, q = 0, g = 0;
    if (c < N) {
        if constexpr (FUSED) g = (double)reinterpret_cast<const float2*>(partial + N + c)->x;       // chunk 0 always has valid rows
#pragma unroll 4
        for (int k = w; k < chunks; k += FIN_WARPS) {
            if constexpr (FUSED) {
                const float2 ss = *reinterpret_cast<const float2*>(partial + (size_t)k * 2 * N + c);
                const float2 sn = *reinterpret_cast<const float2*>(partial + (size_t)k * 2 * N + N + c);
                const double d = (double)sn.x - g, n = (double)sn.y;
                s += (double)ss.x + n * d;
                q += (double)ss.y + 2.0 * d * (double)ss.x + n * d * d;
            } else {
                s += partial[(size_t)k * 2 * N + c];
                q += partial[(size_t)k * 2 * N + N + c];
            }
        }
    }
    sh[w][0][lane] = s; sh[w][1][lane] = q;
    __syncthreads();
    if (w == 0 && c < N) {
        double S = 0, Q = 0;
        for (int k = 0; k < FIN_WARPS; ++k) { S += sh[k][0][lane]; Q += sh[k][1][lane]; }
        const double mean_g = S / M;                      // mean relative to g (g = 0 for the plain sums)
        const double mean = g + mean_g;
        double var = Q / M - mean_g * mean_g;
        if (var < 0) var = 0;
        const double sc = (double)gamma[c] / sqrt(var + (double)eps);
        scale_shift[c] = (float)sc;
        scale_shift[N + c] = (float)((double)beta[c] - mean * sc);
        if (stat_mean) stat_mean[c] = (float)mean;
        if (stat_var) stat_var[c] = (float)var;
    }
}

__device__ __forceinline__ float bn_act(float v, int act) {
    if (act == ACT_RELU) return v < 0.0f ? 0.0f : v;          // like torch.relu: a NaN stays a NaN (fmaxf would return 0)
    if (act == ACT_TANH) return tanhf(v);
    return v;
}

// y[m, c] = act(z[m, c]*scale[c] + shift[c]); 8 channels per thread
template <typename TO>
__global__ void __launch_bounds__(BN_THREADS) bn_apply_kernel(const float* __restrict__ Z, const float* __restrict__ scale_shift,
                                                              TO* __restrict__ out, long long n8, int N, int act) {
    const long long i = (long long)blockIdx.x * BN_THREADS + threadIdx.x;
    if (i >= n8) return;
    const int c = (int)((i * 8) % N);
    const float4 a = *reinterpret_cast<const float4*>(Z + i * 8), b = *reinterpret_cast<const float4*>(Z + i * 8 + 4);
    const float4 s0 = *reinterpret_cast<const float4*>(scale_shift + c), s1 = *reinterpret_cast<const float4*>(scale_shift + c + 4);
    const float4 t0 = *reinterpret_cast<const float4*>(scale_shift + N + c), t1 = *reinterpret_cast<const float4*>(scale_shift + N + c + 4);
    float v[8] = {bn_act(fmaf(a.x, s0.x, t0.x), act), bn_act(fmaf(a.y, s0.y, t0.y), act), bn_act(fmaf(a.z, s0.z, t0.z), act), bn_act(fmaf(a.w, s0.w, t0.w), act),
                  bn_act(fmaf(b.x, s1.x, t1.x), act), bn_act(fmaf(b.y, s1.y, t1.y), act), bn_act(fmaf(b.z, s1.z, t1.z), act), bn_act(fmaf(b.w, s1.w, t1.w), act)};
    if constexpr (sizeof(TO) == 4) {
        *reinterpret_cast<float4*>(out + i * 8) = make_float4(v[0], v[1], v[2], v[3]);
        *reinterpret_cast<float4*>(out + i * 8 + 4) = make_float4(v[4], v[5], v[6], v[7]);
    } else {
        alignas(16) TO o[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) o[j] = TO(v[j]);
        *reinterpret_cast<uint4*>(out + i * 8) = *reinterpret_cast<const uint4*>(o);
    }
}

// out_t[b, c, t] = z[b*T + t, c]*scale[c] + shift[c] (+ res_t[b, c, t]): rows in, the reference's (B,N,T) out
__global__ void __launch_bounds__(BN_THREADS) bn_apply_t_kernel(const float* __restrict__ Z, const float* __restrict__ scale_shift,
                                                                float* __restrict__ out_t, const float* __restrict__ res_t, int T, int N) {
    extern __shared__ float tile[];                      // [32][N + 1]
    const int b = blockIdx.y, t0 = blockIdx.x * 32;
    const int nt = min(32, T - t0);
    for (int i = threadIdx.x; i < nt * N; i += BN_THREADS) {
        const int r = i / N, c = i - r * N;
        tile[r * (N + 1) + c] = fmaf(Z[((size_t)b * T + t0 + r) * N + c], scale_shift[c], scale_shift[N + c]);
    }
    __syncthreads();
    for (int i = threadIdx.x; i < 32 * N; i += BN_THREADS) {
        const int c = i >> 5, r = i & 31;
        if (r >= nt) continue;
        const size_t o = ((size_t)b * N + c) * T + t0 + r;
        out_t[o] = tile[r * (N + 1) + c] + (res_t ? res_t[o] : 0.0f);
    }
}

}  // namespace

int bn_chunks(int M) {
    const int chunks = ceil_div(M, 32);
    return chunks < BN_MAX_CHUNKS ? chunks : BN_MAX_CHUNKS;
}

int bn_partial_capacity(int M) {
    // the GEMM-fused layout has one chunk per (128-row tile, epilogue warp); tiles may be padded (ragged T), at most 2x
    const int fused = 8 * ceil_div(M, 128) + 8;
    const int plain = bn_chunks(M);
    return fused > plain ? fused : plain;
}

cudaError_t launch_bn_train(const BnTrainParams& p, cudaStream_t s, int* launches, int fused_chunks) {
    if (p.N % 8 || p.N > 1024 || p.M <= 0) return cudaErrorInvalidValue;
    const int chunks = fused_chunks > 0 ? fused_chunks : bn_chunks(p.M);
    if (fused_chunks <= 0) {
        const int rows_per_chunk = ceil_div(p.M, chunks);
        const int rpp = BN_THREADS / (p.N / 4);
        bn_partial_kernel<<<chunks, BN_THREADS, (size_t)rpp * 2 * p.N * sizeof(double), s>>>(p.Z, p.partial, p.M, p.N, rows_per_chunk);
    }
    if (fused_chunks > 0) bn_finalize_kernel<true><<<ceil_div(p.N, 32), FIN_WARPS * 32, 0, s>>>(p.partial, chunks, p.M, p.N, p.gamma, p.beta, p.eps, p.scale_shift, p.stat_mean, p.stat_var);
    else bn_finalize_kernel<false><<<ceil_div(p.N, 32), FIN_WARPS * 32, 0, s>>>(p.partial, chunks, p.M, p.N, p.gamma, p.beta, p.eps, p.scale_shift, p.stat_mean, p.stat_var);
    if (p.out_t) {
        dim3 grid(ceil_div(p.T, 32), p.M / p.T);
        bn_apply_t_kernel<<<grid, BN_THREADS, (size_t)32 * (p.N + 1) * sizeof(float), s>>>(p.Z, p.scale_shift, p.out_t, p.res_t, p.T, p.N);
    } else {
        const long long n8 = (long long)p.M * p.N / 8;
        const unsigned grid = (unsigned)((n8 + BN_THREADS - 1) / BN_THREADS);
        if (p.out_type == DT_F32) bn_apply_kernel<float><<<grid, BN_THREADS, 0, s>>>(p.Z, p.scale_shift, (float*)p.out, n8, p.N, p.act);
        else if (p.out_type == DT_BF16) bn_apply_kernel<__nv_bfloat16><<<grid, BN_THREADS, 0, s>>>(p.Z, p.scale_shift, (__nv_bfloat16*)p.out, n8, p.N, p.act);
        else bn_apply_kernel<__half><<<grid, BN_THREADS, 0, s>>>(p.Z, p.scale_shift, (__half*)p.out, n8, p.N, p.act);
    }
    if (launches) *launches = fused_chunks > 0 ? 2 : 3;
    return cudaGetLastError();
}

}  // namespace avc
