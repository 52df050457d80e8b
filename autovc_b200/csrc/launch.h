// Host-side launcher declarations shared between the translation units.
#pragma once
#include "common.cuh"

#include <cstdlib>
#include <cstring>

extern char** environ;

namespace avc {

// Launch of a persistent kernel whose CTAs wait on one another (clusters of `cluster_x` CTAs): cooperative by default, so the
// driver guarantees that the whole grid is co-resident (gang-scheduled).  Profilers that replay kernels (ncu) cannot take a
// cooperative cluster launch; then - and only after cudaOccupancyMaxActiveClusters has confirmed that the grid fits the
// machine - the launch is repeated without the attribute.  `*NOCOOP` environment variables / a detected profiler injection
// skip the cooperative attempt.  reference: none (launch mechanics of this library).
inline bool profiler_attached() {
    static const int v = [] {
        for (char** e = environ; e && *e; ++e)
            if (!strncmp(*e, "NV_COMPUTE_PROFILER_", 20) || !strncmp(*e, "NV_NSIGHT_INJECTION", 19) || !strncmp(*e, "CUDA_INJECTION64_PATH=", 22))
                return 1;
        return 0;
    }();
    return v != 0;
}
inline cudaError_t launch_persistent(const void* fn, dim3 grid, dim3 block, size_t smem, unsigned cluster_x, void** kargs,
                                     cudaStream_t s, bool force_nocoop) {
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = s;
    cudaLaunchAttribute at[2];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = cluster_x; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
    at[1].id = cudaLaunchAttributeCooperative; at[1].val.cooperative = 1;
    cfg.attrs = at;
    if (!force_nocoop && !profiler_attached()) {
        cfg.numAttrs = 2;
        cudaError_t e = cudaLaunchKernelExC(&cfg, fn, kargs);
        if (e == cudaSuccess) return e;
        (void)cudaGetLastError();      // clear the launch error and fall through to the checked non-cooperative launch
    }
    cfg.numAttrs = 1;
    int clusters = 0;
    cudaError_t e = cudaOccupancyMaxActiveClusters(&clusters, fn, &cfg);
    if (e != cudaSuccess) return e;
    const long long need = (long long)grid.x * grid.y * grid.z / (cluster_x ? cluster_x : 1);
    if (clusters < need) return cudaErrorCooperativeLaunchTooLarge;
    return cudaLaunchKernelExC(&cfg, fn, kargs);
}

// simt_kernels.cu
cudaError_t launch_simt_gemm(const GemmParams& p, cudaStream_t s);
cudaError_t launch_simt_lstm_step(const LstmStepParams& q, int ndir, cudaStream_t s);
cudaError_t launch_pack_input(const float* x, const float* c_org, void* X0, int out_type, int B, int T, int E,
                              cudaStream_t s);
// fused-concat path: x (B,80,T) fp32 -> X0 [B][T][80] bf16 and c_org [B][E] fp32 -> Cb bf16, one launch (E even)
cudaError_t launch_pack_mel(const float* x, const float* c_org, void* X0_bf16, void* Cb_bf16, int B, int T, int E, cudaStream_t s);
// encoder conv0 speaker term: S [B][5][N] (per-tap products) + bias [N] -> E [B][5][N] (bias vectors per edge class)
cudaError_t launch_speaker_edge(const float* S, const float* bias, float* E, int B, int N, cudaStream_t s);
cudaError_t launch_codes(const float* Y, const float* c_trg, float* codes, void* D, int d_type, int B, int T, int dn,
                         int E, int freq, cudaStream_t s);

// speaker-encoder helpers (simt_kernels.cu)
cudaError_t launch_cast_pad_rows(const float* in, void* out, int out_type, long long rows, int K, int Kp, cudaStream_t s);
cudaError_t launch_l2norm_rows(const float* in, float* out, int rows, int D, cudaStream_t s);
cudaError_t launch_segment_mean_renorm(const float* emb, const int* offsets, float* out, int S, int D, cudaStream_t s);

// batchnorm.cu (train-mode BatchNorm1d forward on fp32 conv rows)
constexpr int BN_MAX_CHUNKS = 1024;
struct BnTrainParams {
    const float* Z;          // [M][N] fp32 rows: conv output + bias
    int M, N, T;             // T: frames per utterance (channel-first output only)
    const float* gamma;      // [N] BatchNorm weight
    const float* beta;       // [N] BatchNorm bias
    float eps;
    double* partial;         // scratch [bn_chunks(M)][2][N]
    float* scale_shift;      // scratch [2][N]
    float* stat_mean;        // optional out [N]: batch mean
    float* stat_var;         // optional out [N]: biased batch variance
    int act;
    void* out; int out_type; // row-major [M][N] output, or
    float* out_t;            // channel-first fp32 [M/T][N][T] output (+ res_t)
    const float* res_t;
};
int bn_chunks(int M);
int bn_partial_capacity(int M);     // chunks to allocate for BnTrainParams::partial (covers the GEMM-fused layout as well)
// fused_chunks > 0: `partial` already holds that many chunks written by the GEMM epilogue; the statistics pass is skipped
cudaError_t launch_bn_train(const BnTrainParams& p, cudaStream_t s, int* launches, int fused_chunks = 0);

// tc_gemm.cu (tcgen05 + TMA implicit-GEMM)
struct TcGemmPlan;  // opaque: tensor maps + tiling for one (layer, B, T) instance
// Returns nullptr and sets *err when the shape is not supported by the tensor-core kernel.
TcGemmPlan* tc_gemm_plan_create(const GemmParams& p, int num_sms, const char** err);
void tc_gemm_plan_destroy(TcGemmPlan* plan);
// > 0 when the plan's kernel also writes the BatchNorm column sums (GemmParams::stat_partial): that many chunks
int tc_gemm_plan_stat_chunks(const TcGemmPlan* plan);
// out_t/res_t are the caller-owned channel-first output / residual of THIS call (never cached in the plan)
cudaError_t launch_tc_gemm(const TcGemmPlan* plan, cudaStream_t s, float* out_t, const float* res_t);
// one-time driver entry point lookup; returns nullptr on success or an error string
const char* tc_init();

// ll_lstm.cu (persistent weight-stationary LSTM recurrence, tcgen05; h exchanged through L2 with in-band validity)
struct LlLstmPlan;
const char* ll_lstm_init();
size_t ll_lstm_exchange_bytes(int B, int T, int H);
LlLstmPlan* ll_lstm_plan_create(const void* Whh_bf16, const void* Gx, int gx_type, int Tg, int gx_div, void* Y, void* Yx, int B, int T, int H,
                                int num_sms, const char** err);
void ll_lstm_plan_destroy(LlLstmPlan* plan);
cudaError_t launch_ll_lstm(const LlLstmPlan* plan, cudaStream_t s, int* launches);

// enc_lstm.cu (encoder BiLSTM, H=32: one warp per 16 utterances x direction, W_hh in registers, mma.sync)
bool enc_lstm_supported(int dim_neck);
cudaError_t launch_enc_lstm_l0(const float* Gx, const void* Whh_bf16, void* Y_bf16, int B, int T, cudaStream_t s);
cudaError_t launch_enc_lstm_l1(const float* Gx, const void* Whh_bf16, float* codes, void* D_bf16, int ldd, const float* c_trg,
                               int E, int B, int T, int freq, cudaStream_t s, int* launches);

}  // namespace avc
