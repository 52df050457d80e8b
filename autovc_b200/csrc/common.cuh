// Shared definitions for the autovc_b200 CUDA sources (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <stdint.h>

namespace avc {

// element types of activation / weight buffers
enum DType : int { DT_F32 = 0, DT_BF16 = 1, DT_F16 = 2 };
__host__ __device__ inline int dtype_size(int dt) { return dt == DT_F32 ? 4 : 2; }

enum Act : int { ACT_NONE = 0, ACT_RELU = 1, ACT_TANH = 2 };

constexpr int N_MELS = 80;
constexpr int CONV_CH = 512;     // encoder / postnet conv width (reference encoder.py:41, postnet.py:23)
constexpr int LSTM2_H = 1024;    // reference decoder.py:53
constexpr int KTAPS = 5;         // reference encoder.py:42 / decoder.py:44 / postnet.py:24 (kernel_size=5, padding=2)

// ---------------------------------------------------------------------------------------------
// A GEMM-shaped layer:  out[m,n] = act( sum_k A[m,k] * W[n,k] + bias[n] )
// Convolutions are implicit GEMMs: k = tap*Cp + ci, A[m,k] = in[b, t+tap-taps/2, ci] (0 outside
// the utterance or for ci >= Cin), m = b*T + t.  Plain GEMMs use taps = 1.
// ---------------------------------------------------------------------------------------------
struct GemmParams {
    const void* A;      // activation rows, [rows][lda] elements of a_type
    int a_type;
    int lda;
    int M, N, K;        // K = taps * Cp
    int T;              // rows per utterance (halo logic); plain GEMM: any value, taps = 1
    int taps, Cin, Cp;
    const void* W;      // packed weights [N][ldw] of w_type
    int w_type;
    int ldw;
    const float* bias;  // [N]
    const float* row_bias;  // optional (conv layers, tensor-core path): [B][5][N] added per row instead of `bias`; class of frame t:
                            // 0 interior, 1: t = 0, 2: t = 1, 3: t = T-2, 4: t = T-1  (the fused speaker term of encoder conv0)
    int act;
    void* out;          // row-major [M][ldo] of out_type, or nullptr
    int out_type;
    int ldo;
    float* out_t;       // optional channel-first fp32 output [M/T][N][T] (the reference's (B,80,T) layout)
    const float* res_t; // optional residual added to out_t (same layout)
    double* stat_partial;   // optional (tensor-core path, fp32 rows): fused per-column sums for the train-mode BatchNorm, [chunk][2][N]
    int stat_capacity;      // chunks the buffer holds
};

// One time step of an LSTM layer (both directions in one launch, blockIdx.z = direction):
//   gates[b, u*4+g] = Gx[b, u*4+g] + sum_k h_prev[b,k] * Whh[u*4+g, k]        (g: 0=i 1=f 2=g 3=o)
//   c' = sig(f) c + sig(i) tanh(g);  h' = sig(o) tanh(c')       (PyTorch nn.LSTM; reference encoder.py:51)
struct LstmStepParams {
    const void* h_prev[2];  // [B] rows, pitch ldh elements, of h_type; nullptr -> zero initial state
    int h_type;
    int ldh;
    const void* Whh[2];     // packed [4H][H] (unit-major gate rows) of w_type
    int w_type;
    const float* Gx[2];     // row b at Gx + b*ldg, 4H contiguous floats (unit-major), bias included
    long long ldg;
    float* c[2];            // [B][H] fp32 cell state, updated in place
    void* h_out[2];         // row b at h_out + b*ldho elements (h_type)
    int ldho;
    int B, H;
};

inline __host__ __device__ int ceil_div(int a, int b) { return (a + b - 1) / b; }
inline __host__ __device__ int round_up(int a, int b) { return (a + b - 1) / b * b; }

}  // namespace avc
