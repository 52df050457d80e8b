// C ABI of autovc_b200 (see include/autovc_b200.h): context, weight folding/packing, workspace
// carving and the forward-pass schedule.  Host code only; kernels live in the other .cu files.
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <cmath>
#include <map>
#include <string>
#include <vector>

#include "../../include/autovc_b200.h"
#include "common.cuh"
#include "launch.h"

using namespace avc;

namespace {

thread_local std::string g_create_error = "";

struct Layer {
    std::string name;
    int N = 0, K = 0, Kp = 0;        // logical K, padded K (row pitch)
    int taps = 1, Cin = 0, Cp = 0;   // conv geometry (taps = 1: plain GEMM, Cin = K, Cp = Kp)
    int w_type = DT_F32;
    std::vector<float> w, b;         // fp32 masters of the packed matrix [N][Kp] and bias [N]
    void* w_dev = nullptr;
    float* b_dev = nullptr;
};

enum LayerId {
    L_ENC_CONV0 = 0, L_ENC_CONV1, L_ENC_CONV2,
    L_ENC_L0_IH, L_ENC_L0_HH, L_ENC_L1_IH, L_ENC_L1_HH,
    L_DEC_LSTM1_IH, L_DEC_LSTM1_HH,
    L_DEC_CONV0, L_DEC_CONV1, L_DEC_CONV2,
    L_DEC_L2_0_IH, L_DEC_L2_0_HH, L_DEC_L2_1_IH, L_DEC_L2_1_HH,
    L_DEC_LINEAR,
    L_POST_CONV0, L_POST_CONV1, L_POST_CONV2, L_POST_CONV3, L_POST_CONV4,
    L_GEN_COUNT,
    // d-vector speaker encoder (reference autovc/speaker_encoder/model.py:48-58): LSTM(40,256,3) + Linear(256,256)
    L_SE_L0_IH = L_GEN_COUNT, L_SE_L0_HH, L_SE_L1_IH, L_SE_L1_HH, L_SE_L2_IH, L_SE_L2_HH, L_SE_LINEAR,
    // the 11 convolutions again WITHOUT the BatchNorm folded in, for the train-mode (batch statistics) forward
    L_RAW_CONV0, L_RAW_COUNT_END = L_RAW_CONV0 + 11,
    // encoder conv0 split for the fused speaker concat (reference encoder.py:65-66): the 80 mel channels as a 5-tap convolution,
    // the dim_emb speaker channels as a plain GEMM per tap whose result enters conv0's epilogue as a per-utterance bias
    L_ENC_CONV0_MEL = L_RAW_COUNT_END, L_ENC_CONV0_SPK,
    L_COUNT
};

constexpr int N_BN = 11;   // BatchNorm layers in path order: encoder 0-2, decoder 0-2, postnet 0-4
constexpr int BN_MAX_CH = 512;
const int kBnFolded[N_BN] = {L_ENC_CONV0, L_ENC_CONV1, L_ENC_CONV2, L_DEC_CONV0, L_DEC_CONV1, L_DEC_CONV2,
                             L_POST_CONV0, L_POST_CONV1, L_POST_CONV2, L_POST_CONV3, L_POST_CONV4};

constexpr int SE_MELS = 40, SE_H = 256, SE_LAYERS = 3, SE_EMB = 256;   // reference utils/hparams.py:60-63, model.py:48-58

const char* kLayerNames[L_COUNT] = {
    "enc.conv0", "enc.conv1", "enc.conv2",
    "enc.lstm.l0.ih", "enc.lstm.l0.hh", "enc.lstm.l1.ih", "enc.lstm.l1.hh",
    "dec.lstm1.ih", "dec.lstm1.hh",
    "dec.conv0", "dec.conv1", "dec.conv2",
    "dec.lstm2.l0.ih", "dec.lstm2.l0.hh", "dec.lstm2.l1.ih", "dec.lstm2.l1.hh",
    "dec.linear",
    "post.conv0", "post.conv1", "post.conv2", "post.conv3", "post.conv4",
    "se.lstm.l0.ih", "se.lstm.l0.hh", "se.lstm.l1.ih", "se.lstm.l1.hh", "se.lstm.l2.ih", "se.lstm.l2.hh", "se.linear",
    "raw.enc.conv0", "raw.enc.conv1", "raw.enc.conv2", "raw.dec.conv0", "raw.dec.conv1", "raw.dec.conv2",
    "raw.post.conv0", "raw.post.conv1", "raw.post.conv2", "raw.post.conv3", "raw.post.conv4",
    "enc.conv0.mel", "enc.conv0.spk",
};

struct Workspace {
    // all pointers are carved from the caller's workspace; sizes in bytes
    char* X0; char* actA; char* actB; float* GxE; char* Y0; float* Y1; float* cstate;
    char* Dm; float* Gd1; char* H1; float* G2; char* H2a; char* H2b; char* melc;
    char* Yb;    // exchange copy of the running LSTM layer's output (ll_lstm.cu)
    char* Cb; float* Sspk; float* Espk;   // fused speaker concat: c_org in bf16 [B][E], per-tap products [B][5][512], edge-class bias vectors [B][5][512]
    double* bn_partial; float* bn_ss;   // train-mode BatchNorm scratch (batchnorm.cu); the fp32 conv rows reuse G2
    size_t total;
};

// mode: 0 forward, 1 encode only, 2 speaker encoder, 3 / 4 forward / encode with batch-statistics BatchNorm; + 8: conv0 over the materialised concat
struct PlanSet {
    void* ws = nullptr; int B = 0, T = 0, mode = 0;
    uint64_t stamp = 0;
    std::vector<TcGemmPlan*> plans;
    std::vector<LlLstmPlan*> llplans;
    // CUDA graphs of this plan set's whole launch sequence, keyed on the caller's pointers (see run()): at most GRAPHS_PER_SET,
    // least recently used first out
    struct GraphEntry {
        const void* key[7] = {};       // x, c_org, c_trg, mel, mel_post, codes, bn_stats
        cudaGraphExec_t exec = nullptr;
        int seen = 0;                  // calls with this key so far (the graph is built on the second)
        bool failed = false;           // capture or instantiation failed once: this key keeps the plain launch sequence
        int64_t launches = 0;
        uint64_t stamp = 0;
    };
    std::vector<GraphEntry> graphs;
    int captures = 0;              // graphs built for this plan set ...
    int64_t replays = 0;           // ... and launches they served: building is rationed by their ratio (see run())
    ~PlanSet() {
        for (auto* p : plans) tc_gemm_plan_destroy(p);
        for (auto* p : llplans) ll_lstm_plan_destroy(p);
        for (auto& g : graphs) if (g.exec) cudaGraphExecDestroy(g.exec);
    }
};
constexpr size_t GRAPHS_PER_SET = 6;
constexpr size_t PLAN_CACHE_ENTRIES = 8;

// restores the caller's current device when an entry point returns (the library switches to the context's device)
struct DeviceGuard {
    int prev = -1;
    cudaError_t enter(int device) {
        if (cudaGetDevice(&prev) != cudaSuccess) prev = -1;
        return prev == device ? cudaSuccess : cudaSetDevice(device);
    }
    ~DeviceGuard() { int cur = -1; if (prev >= 0 && cudaGetDevice(&cur) == cudaSuccess && cur != prev) cudaSetDevice(prev); }
};

}  // namespace

struct avc_ctx {
    avc_dims dims{};
    int device = -1;
    int precision = AVC_PREC_BF16;
    int num_sms = 148;
    bool loaded = false;
    bool gx_f32 = false;       // AVC_GX_F32: keep lstm2's gate pre-activations in fp32 (A/B aid)
    bool se_loaded = false;    // avc_se_load_weights completed
    bool force_simt = false;   // AVC_FORCE_SIMT=1: debugging aid, run the GEMMs of the bf16 mode on the CUDA cores
    std::string err = "";
    std::map<std::string, std::vector<float>> raw;
    Layer layers[L_COUNT];
    float* bn_gamma_dev[N_BN] = {};   // BatchNorm weight / bias of the 11 conv blocks (train-mode forward)
    float* bn_beta_dev[N_BN] = {};
    int64_t launches = 0;
    // tensor-core plan cache: a small LRU of plan sets (tensor maps + launch geometry), one per (workspace, B, T, mode).
    // Alternating shapes - the converter's length buckets, a B=1 call after a B=256 step, the training pair's
    // encode-only pass - hit their own entry instead of re-encoding every tensor map.
    std::vector<PlanSet*> plan_cache;
    uint64_t plan_clock = 0;
    int64_t plan_builds = 0;
    int64_t graph_replays = 0;   // forwards that ran as one cudaGraphLaunch
    bool no_graph = false;       // AVC_NO_GRAPH=1 / AVC_LSTM_TRACE: always the plain launch sequence
    cudaStream_t capture_stream = nullptr;   // private stream the graphs are captured on
    bool simt_lstm = false;    // AVC_SIMT_LSTM=1: debugging aid, per-step CUDA-core recurrences in the bf16 mode
    // optional per-phase timing (avc_set_profiling): events at phase boundaries of the last run
    bool profiling = false;
    std::vector<cudaEvent_t> prof_events;
    std::vector<int> prof_cats;
};

namespace {

int fail(avc_ctx* h, int code, const char* fmt, ...) {
    char buf[1024];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    if (h) h->err = buf; else g_create_error = buf;
    return code;
}

// unidirectional LSTM layers of this hidden size run on the tensor-core recurrence (ll_lstm.cu): decides the kernel at run
// time AND the gate-row prescale at pack time, so both must read the same predicate
bool ll_eligible(const avc_ctx* h, int H) {
    return h->precision == AVC_PREC_BF16 && !h->force_simt && !h->simt_lstm && H % 256 == 0 && H <= 1024;
}

int act_type(const avc_ctx* h) { return h->precision == AVC_PREC_FP32 ? DT_F32 : DT_BF16; }
int post_type(const avc_ctx* h) { return h->precision == AVC_PREC_FP32 ? DT_F32 : DT_F16; }

size_t align256(size_t v) { return (v + 255) & ~size_t(255); }

// ---------------------------------------------------------------------------------------------
// weight folding / packing (host)
// ---------------------------------------------------------------------------------------------
const std::vector<float>* find_raw(avc_ctx* h, const std::string& name, int64_t numel) {
    auto it = h->raw.find(name);
    if (it == h->raw.end()) {
        fail(h, AVC_ERR_MISSING_WEIGHT, "state-dict entry '%s' is missing", name.c_str());
        return nullptr;
    }
    if ((int64_t)it->second.size() != numel) {
        fail(h, AVC_ERR_MISSING_WEIGHT, "state-dict entry '%s' has %lld elements, expected %lld", name.c_str(),
             (long long)it->second.size(), (long long)numel);
        return nullptr;
    }
    return &it->second;
}

// Conv1d(k=5) + eval BatchNorm -> packed [Cout][5*Cp], k = tap*Cp + ci, and folded bias.
// reference net_layers.py:39-49 (conv), encoder.py:46 (BatchNorm1d, eps 1e-5).
bool pack_conv(avc_ctx* h, Layer& L, const std::string& prefix, int cin, int cout, bool fold = true) {
    const auto* w = find_raw(h, prefix + ".0.conv.weight", (int64_t)cout * cin * KTAPS);
    const auto* b = find_raw(h, prefix + ".0.conv.bias", cout);
    const auto* g = find_raw(h, prefix + ".1.weight", cout);
    const auto* be = find_raw(h, prefix + ".1.bias", cout);
    const auto* mu = find_raw(h, prefix + ".1.running_mean", cout);
    const auto* var = find_raw(h, prefix + ".1.running_var", cout);
    if (!w || !b || !g || !be || !mu || !var) return false;
    L.taps = KTAPS; L.Cin = cin; L.Cp = round_up(cin, 64);
    L.N = cout; L.K = KTAPS * L.Cp; L.Kp = L.K;
    L.w.assign((size_t)L.N * L.Kp, 0.0f);
    L.b.assign(L.N, 0.0f);
    for (int n = 0; n < cout; ++n) {
        const double s = fold ? (double)(*g)[n] / std::sqrt((double)(*var)[n] + 1e-5) : 1.0;
        L.b[n] = fold ? (float)(((double)(*b)[n] - (double)(*mu)[n]) * s + (double)(*be)[n]) : (*b)[n];
        for (int ci = 0; ci < cin; ++ci)
            for (int tap = 0; tap < KTAPS; ++tap)
                L.w[(size_t)n * L.Kp + tap * L.Cp + ci] = (float)((double)(*w)[((size_t)n * cin + ci) * KTAPS + tap] * s);
    }
    return true;
}

// Encoder conv0 split along the input channels (fused speaker concat): `mel` = channels [0, 80) as a 5-tap convolution
// [Cout][5*128], `spk` = channels [80, 80+E) as a plain matrix [5*Cout][E] (row tap*Cout + n), both with the BatchNorm scale
// folded in; the folded bias stays with the edge-class vectors (speaker_edge_kernel).  Built from the packed full layer.
void split_conv0(const Layer& F, Layer& M, Layer& S, int n_mels, int E) {
    M.taps = KTAPS; M.Cin = n_mels; M.Cp = round_up(n_mels, 64); M.N = F.N; M.K = KTAPS * M.Cp; M.Kp = M.K;
    M.w.assign((size_t)M.N * M.Kp, 0.0f); M.b = F.b;
    S.taps = 1; S.Cin = E; S.Cp = E; S.N = KTAPS * F.N; S.K = E; S.Kp = E;
    S.w.assign((size_t)S.N * S.Kp, 0.0f); S.b.assign(S.N, 0.0f);
    for (int n = 0; n < F.N; ++n)
        for (int tap = 0; tap < KTAPS; ++tap) {
            const float* src = &F.w[(size_t)n * F.Kp + tap * F.Cp];
            for (int c = 0; c < n_mels; ++c) M.w[(size_t)n * M.Kp + tap * M.Cp + c] = src[c];
            for (int c = 0; c < E; ++c) S.w[(size_t)(tap * F.N + n) * S.Kp + c] = src[n_mels + c];
        }
}

// nn.LSTM matrices -> unit-major gate rows: packed row d*4H + u*4 + g  <-  PyTorch row g*H + u of direction d.
// ih: [ndir*4H][K] with bias = b_ih + b_hh;  hh: [ndir*4H][H], no bias.
// `prescale` (layers that run on the tensor-core recurrence, ll_lstm.cu): the rows of the sigmoid gates (i, f, o) of W_ih, W_hh
// and the bias are multiplied by 0.5 - exact in every binary format - so that the kernel's gate pre-activation already is the
// argument of  sigmoid(x) = 0.5 tanh(0.5 x) + 0.5  and the epilogue saves one multiply per gate value.
bool pack_lstm(avc_ctx* h, Layer& Lih, Layer& Lhh, const std::string& prefix, int layer, int in_k, int H, bool bidir, bool prescale = false) {
    const int ndir = bidir ? 2 : 1;
    Lih.taps = 1; Lih.N = ndir * 4 * H; Lih.K = in_k; Lih.Kp = round_up(in_k, 16); Lih.Cin = in_k; Lih.Cp = Lih.Kp;
    Lhh.taps = 1; Lhh.N = ndir * 4 * H; Lhh.K = H; Lhh.Kp = H; Lhh.Cin = H; Lhh.Cp = H;
    Lih.w.assign((size_t)Lih.N * Lih.Kp, 0.0f); Lih.b.assign(Lih.N, 0.0f);
    Lhh.w.assign((size_t)Lhh.N * Lhh.Kp, 0.0f); Lhh.b.assign(Lhh.N, 0.0f);
    for (int d = 0; d < ndir; ++d) {
        const std::string sfx = "_l" + std::to_string(layer) + (d ? "_reverse" : "");
        const auto* wih = find_raw(h, prefix + ".weight_ih" + sfx, (int64_t)4 * H * in_k);
        const auto* whh = find_raw(h, prefix + ".weight_hh" + sfx, (int64_t)4 * H * H);
        const auto* bih = find_raw(h, prefix + ".bias_ih" + sfx, 4 * H);
        const auto* bhh = find_raw(h, prefix + ".bias_hh" + sfx, 4 * H);
        if (!wih || !whh || !bih || !bhh) return false;
        for (int u = 0; u < H; ++u)
            for (int g = 0; g < 4; ++g) {
                const int src = g * H + u, dst = d * 4 * H + u * 4 + g;
                memcpy(&Lih.w[(size_t)dst * Lih.Kp], &(*wih)[(size_t)src * in_k], sizeof(float) * in_k);
                memcpy(&Lhh.w[(size_t)dst * Lhh.Kp], &(*whh)[(size_t)src * H], sizeof(float) * H);
                Lih.b[dst] = (*bih)[src] + (*bhh)[src];
                if (prescale && g != 2) {
                    for (int k = 0; k < in_k; ++k) Lih.w[(size_t)dst * Lih.Kp + k] *= 0.5f;
                    for (int k = 0; k < H; ++k) Lhh.w[(size_t)dst * Lhh.Kp + k] *= 0.5f;
                    Lih.b[dst] *= 0.5f;
                }
            }
    }
    return true;
}

bool pack_linear(avc_ctx* h, Layer& L, const std::string& prefix, int in_k, int out_n) {
    const auto* w = find_raw(h, prefix + ".weight", (int64_t)out_n * in_k);
    const auto* b = find_raw(h, prefix + ".bias", out_n);
    if (!w || !b) return false;
    L.taps = 1; L.N = out_n; L.K = in_k; L.Kp = round_up(in_k, 16); L.Cin = in_k; L.Cp = L.Kp;
    L.w.assign((size_t)L.N * L.Kp, 0.0f);
    for (int n = 0; n < out_n; ++n) memcpy(&L.w[(size_t)n * L.Kp], &(*w)[(size_t)n * in_k], sizeof(float) * in_k);
    L.b = *b;
    return true;
}

int upload_layer(avc_ctx* h, Layer& L) {
    const size_t n = L.w.size();
    cudaError_t e;
    if (L.w_dev) { cudaFree(L.w_dev); L.w_dev = nullptr; }
    if (L.b_dev) { cudaFree(L.b_dev); L.b_dev = nullptr; }
    if (L.w_type == DT_F32) {
        if ((e = cudaMalloc(&L.w_dev, n * 4)) != cudaSuccess) goto bad;
        if ((e = cudaMemcpy(L.w_dev, L.w.data(), n * 4, cudaMemcpyHostToDevice)) != cudaSuccess) goto bad;
    } else {
        std::vector<uint16_t> tmp(n);
        if (L.w_type == DT_BF16) {
            for (size_t i = 0; i < n; ++i) { __nv_bfloat16 v = __float2bfloat16_rn(L.w[i]); memcpy(&tmp[i], &v, 2); }
        } else {
            for (size_t i = 0; i < n; ++i) {
                if (std::fabs(L.w[i]) > 65504.0f)
                    return fail(h, AVC_ERR_UNSUPPORTED, "layer %s: folded weight %g exceeds the fp16 range; use AVC_PREC_FP32",
                                L.name.c_str(), (double)L.w[i]);
                __half v = __float2half_rn(L.w[i]); memcpy(&tmp[i], &v, 2);
            }
        }
        if ((e = cudaMalloc(&L.w_dev, n * 2)) != cudaSuccess) goto bad;
        if ((e = cudaMemcpy(L.w_dev, tmp.data(), n * 2, cudaMemcpyHostToDevice)) != cudaSuccess) goto bad;
    }
    if ((e = cudaMalloc((void**)&L.b_dev, L.b.size() * 4)) != cudaSuccess) goto bad;
    if ((e = cudaMemcpy(L.b_dev, L.b.data(), L.b.size() * 4, cudaMemcpyHostToDevice)) != cudaSuccess) goto bad;
    return AVC_OK;
bad:
    return fail(h, AVC_ERR_CUDA, "uploading layer %s: %s", L.name.c_str(), cudaGetErrorString(e));
}

// ---------------------------------------------------------------------------------------------
// workspace
// ---------------------------------------------------------------------------------------------
Workspace carve(const avc_ctx* h, char* base, int B, int T) {
    const avc_dims& d = h->dims;
    const size_t as = dtype_size(act_type(h));
    const size_t M = (size_t)B * T;
    const int nb = ceil_div(T, d.freq);
    const int cw = d.dim_pre > CONV_CH ? d.dim_pre : CONV_CH;
    const int hmax = d.dim_pre > LSTM2_H ? d.dim_pre : LSTM2_H;
    Workspace w{};
    size_t off = 0;
    auto take = [&](size_t bytes) { char* p = base ? base + off : nullptr; off += align256(bytes); return p; };
    w.X0 = take(M * (N_MELS + d.dim_emb) * as);
    w.actA = take(M * cw * as);
    w.actB = take(M * cw * as);
    w.GxE = (float*)take(M * 8 * d.dim_neck * 4);
    w.Y0 = take(M * 2 * d.dim_neck * as);
    w.Y1 = (float*)take(M * 2 * d.dim_neck * 4);
    w.cstate = (float*)take((size_t)2 * B * hmax * 4);
    w.Dm = take((size_t)B * nb * (2 * d.dim_neck + d.dim_emb) * as);
    w.Gd1 = (float*)take((size_t)B * nb * 4 * d.dim_pre * 4);
    w.H1 = take(M * d.dim_pre * as);
    w.G2 = (float*)take(M * 4 * LSTM2_H * 4);
    w.H2a = take(M * LSTM2_H * as);
    w.H2b = take(M * LSTM2_H * as);
    w.melc = take(M * N_MELS * as);
    w.Yb = take(ll_lstm_exchange_bytes(B, T, hmax));
    w.Cb = take((size_t)B * d.dim_emb * 2);
    w.Sspk = (float*)take((size_t)B * KTAPS * CONV_CH * 4);
    w.Espk = (float*)take((size_t)B * KTAPS * CONV_CH * 4);
    w.bn_partial = (double*)take((size_t)bn_partial_capacity((int)M) * 2 * BN_MAX_CH * sizeof(double));
    w.bn_ss = (float*)take((size_t)2 * BN_MAX_CH * sizeof(float));
    w.total = off;
    return w;
}

// ---------------------------------------------------------------------------------------------
// schedule
// ---------------------------------------------------------------------------------------------
struct Runner {
    avc_ctx* h;
    cudaStream_t s;
    int B, T;
    int rc = AVC_OK;
    PlanSet* ps = nullptr;      // the plan set of this (workspace, B, T, mode); filled on its first run
    size_t plan_cursor = 0, llplan_cursor = 0;
    bool building_plans = false;

    bool ok() const { return rc == AVC_OK; }
    void check(cudaError_t e, const char* what) {
        if (rc == AVC_OK && e != cudaSuccess) rc = fail(h, AVC_ERR_CUDA, "%s: %s", what, cudaGetErrorString(e));
    }

    bool use_tc() const { return h->precision == AVC_PREC_BF16 && !h->force_simt; }

    // phase boundary marker for avc_get_profile (no-op unless profiling is on)
    int cur_cat = -1;
    void mark(int cat) {
        if (!h->profiling || cat == cur_cat || !ok()) return;
        cudaEvent_t ev;
        if (cudaEventCreate(&ev) != cudaSuccess) return;
        cudaEventRecord(ev, s);
        h->prof_events.push_back(ev);
        h->prof_cats.push_back(cat);
        cur_cat = cat;
    }

    // out[m,n] = act(A W^T + b); A is an activation buffer of `a_type`
    int last_stat_chunks = 0;   // set by gemm(): > 0 when the launch also wrote the BatchNorm column sums
    void gemm(int lid, const void* A, int a_type, int lda, int M, int Tutt, int act, void* out, int out_type, int ldo,
              float* out_t = nullptr, const float* res_t = nullptr, double* stat_partial = nullptr, const float* row_bias = nullptr) {
        if (!ok()) return;
        const Layer& L = h->layers[lid];
        mark(L.taps > 1 ? AVC_PHASE_CONV : AVC_PHASE_PROJ);
        GemmParams p{};
        p.A = A; p.a_type = a_type; p.lda = lda;
        p.M = M; p.N = L.N; p.K = L.Kp; p.T = Tutt;
        p.taps = L.taps; p.Cin = L.Cin; p.Cp = L.Cp;
        p.W = L.w_dev; p.w_type = L.w_type; p.ldw = L.Kp;
        p.bias = row_bias ? nullptr : L.b_dev; p.row_bias = row_bias; p.act = act;
        p.out = out; p.out_type = out_type; p.ldo = ldo;
        p.out_t = out_t; p.res_t = res_t;
        p.stat_partial = stat_partial; p.stat_capacity = stat_partial ? bn_partial_capacity(M) : 0;
        last_stat_chunks = 0;
        if (use_tc()) {
            if (building_plans) {
                const char* why = nullptr;
                TcGemmPlan* plan = tc_gemm_plan_create(p, h->num_sms, &why);
                if (!plan) { rc = fail(h, AVC_ERR_UNSUPPORTED, "tensor-core GEMM plan for %s: %s", L.name.c_str(), why ? why : "?"); return; }
                ps->plans.push_back(plan);
            }
            last_stat_chunks = tc_gemm_plan_stat_chunks(ps->plans[plan_cursor]);
            check(launch_tc_gemm(ps->plans[plan_cursor++], s, out_t, res_t), L.name.c_str());
        } else {
            check(launch_simt_gemm(p, s), L.name.c_str());
        }
        h->launches++;
    }

    // conv block `bn` (0..10): eval mode = one GEMM with the folded weights; train mode (bn_train) = GEMM with the unfolded
    // weights -> fp32 rows -> batch statistics -> normalise + activation (batchnorm.cu).  reference encoder.py:69-70 etc.
    bool bn_train = false;
    float* bn_stats = nullptr;   // caller's [N_BN][2][BN_MAX_CH] (mean, biased var) or nullptr
    void conv_bn(int bn, const void* A, int a_type, int lda, int M, int Tutt, int act, void* out, int out_type, int ldo,
                 float* out_t = nullptr, const float* res_t = nullptr) {
        if (!bn_train) { gemm(kBnFolded[bn], A, a_type, lda, M, Tutt, act, out, out_type, ldo, out_t, res_t); return; }
        const Layer& L = h->layers[L_RAW_CONV0 + bn];
        float* Z = wsp->G2;
        gemm(L_RAW_CONV0 + bn, A, a_type, lda, M, Tutt, ACT_NONE, Z, DT_F32, L.N, nullptr, nullptr,
             getenv("AVC_BN_NO_FUSED_STATS") ? nullptr : wsp->bn_partial);
        if (!ok()) return;
        const int fused_chunks = last_stat_chunks;
        BnTrainParams q{};
        q.Z = Z; q.M = M; q.N = L.N; q.T = Tutt;
        q.gamma = h->bn_gamma_dev[bn]; q.beta = h->bn_beta_dev[bn]; q.eps = 1e-5f;
        q.partial = wsp->bn_partial; q.scale_shift = wsp->bn_ss;
        q.stat_mean = bn_stats ? bn_stats + (size_t)bn * 2 * BN_MAX_CH : nullptr;
        q.stat_var = bn_stats ? bn_stats + (size_t)bn * 2 * BN_MAX_CH + BN_MAX_CH : nullptr;
        q.act = act; q.out = out; q.out_type = out_type; q.out_t = out_t; q.res_t = res_t;
        int n = 0;
        check(launch_bn_train(q, s, &n, fused_chunks), "train-mode BatchNorm");
        h->launches += n;
    }

    // Full-sequence LSTM layer from a hoisted input projection Gx.
    //   Gx: [B][Tg][ndir*4H] fp32 (row for frame t is t/gx_div);  Y: [B][T][ndir*H] of h_type.
    // true when lstm() takes the in-band-exchange tensor-core kernel for this layer (which also accepts fp16 Gx)
    bool ll_path(int lid_hh, int h_type, int H, int ndir) const {
        return ll_eligible(h, H) && ndir == 1 && h_type == DT_BF16 && h->layers[lid_hh].w_type == DT_BF16;
    }
    // fp16 gate pre-activations for the big recurrent layers: halves the Gx write + read (1 GB per forward at config 2)
    int gx_type(int lid_hh, int h_type, int H) const { return ll_path(lid_hh, h_type, H, 1) && !h->gx_f32 ? DT_F16 : DT_F32; }

    void lstm(int lid_hh, const void* Gx_, int gx_t, int Tg, int gx_div, void* Y, int h_type, int H, int ndir) {
        const float* Gx = (const float*)Gx_;
        if (!ok()) return;
        const Layer& L = h->layers[lid_hh];
        mark(lid_hh <= L_ENC_L1_HH ? AVC_PHASE_ENC_LSTM : (lid_hh == L_DEC_LSTM1_HH ? AVC_PHASE_LSTM1 : AVC_PHASE_LSTM2));
        const size_t hs = dtype_size(h_type);
        if (ll_path(lid_hh, h_type, H, ndir)) {
            // persistent tensor-core recurrence, in-band validity exchange (ll_lstm.cu)
            if (building_plans) {
                const char* why = nullptr;
                LlLstmPlan* plan = ll_lstm_plan_create(L.w_dev, Gx_, gx_t, Tg, gx_div, Y, wsp->Yb, B, T, H, h->num_sms, &why);
                if (!plan) { rc = fail(h, AVC_ERR_UNSUPPORTED, "tensor-core LSTM plan for %s: %s", L.name.c_str(), why ? why : "?"); return; }
                ps->llplans.push_back(plan);
            }
            int n = 0;
            check(launch_ll_lstm(ps->llplans[llplan_cursor++], s, &n), L.name.c_str());
            h->launches += n;
            return;
        }
        if (gx_t != DT_F32) { rc = fail(h, AVC_ERR_UNSUPPORTED, "fp16 Gx outside the ll_lstm path"); return; }
        check(cudaMemsetAsync(wsp->cstate, 0, (size_t)2 * B * H * 4, s), "memset c");
        const int ldy = ndir * H;
        for (int step = 0; step < T && ok(); ++step) {
            LstmStepParams q{};
            q.h_type = h_type; q.w_type = L.w_type; q.B = B; q.H = H;
            q.ldh = T * ldy; q.ldho = T * ldy; q.ldg = (long long)Tg * ndir * 4 * H;
            for (int d = 0; d < ndir; ++d) {
                const int t = d == 0 ? step : T - 1 - step;
                const int tp = d == 0 ? t - 1 : t + 1;
                q.h_prev[d] = step == 0 ? nullptr : (const char*)Y + ((size_t)tp * ldy + d * H) * hs;
                q.h_out[d] = (char*)Y + ((size_t)t * ldy + d * H) * hs;
                q.Whh[d] = (const char*)L.w_dev + (size_t)d * 4 * H * H * dtype_size(L.w_type);
                q.Gx[d] = Gx + (size_t)(t / gx_div) * ndir * 4 * H + (size_t)d * 4 * H;
                q.c[d] = wsp->cstate + (size_t)d * B * H;
            }
            check(launch_simt_lstm_step(q, ndir, s), L.name.c_str());
            h->launches++;
        }
    }

    Workspace* wsp = nullptr;
};

// ---- plan cache (LRU) ----
void drop_plans(avc_ctx* h, int mode_filter = -1) {      // mode_filter < 0: every entry; 2: the speaker encoder's; 0: the generator's (modes != 2)
    std::vector<PlanSet*> keep;
    for (auto* ps : h->plan_cache) {
        const bool drop = mode_filter < 0 || (mode_filter == 2 ? ps->mode == 2 : ps->mode != 2);
        if (drop) delete ps; else keep.push_back(ps);
    }
    h->plan_cache.swap(keep);
}
void drop_plan_set(avc_ctx* h, PlanSet* ps) {
    for (size_t i = 0; i < h->plan_cache.size(); ++i)
        if (h->plan_cache[i] == ps) { h->plan_cache.erase(h->plan_cache.begin() + i); break; }
    delete ps;
}
// the entry for (workspace, B, T, mode), created empty (least recently used entry evicted) when absent
PlanSet* plan_set_for(avc_ctx* h, void* ws, int B, int T, int mode, bool* fresh) {
    for (auto* ps : h->plan_cache)
        if (ps->ws == ws && ps->B == B && ps->T == T && ps->mode == mode) { ps->stamp = ++h->plan_clock; *fresh = false; return ps; }
    if (h->plan_cache.size() >= PLAN_CACHE_ENTRIES) {
        size_t lru = 0;
        for (size_t i = 1; i < h->plan_cache.size(); ++i)
            if (h->plan_cache[i]->stamp < h->plan_cache[lru]->stamp) lru = i;
        delete h->plan_cache[lru];
        h->plan_cache.erase(h->plan_cache.begin() + lru);
    }
    PlanSet* ps = new PlanSet();
    ps->ws = ws; ps->B = B; ps->T = T; ps->mode = mode; ps->stamp = ++h->plan_clock;
    h->plan_cache.push_back(ps);
    h->plan_builds++;
    *fresh = true;
    return ps;
}

int run(avc_ctx* h, const float* x, const float* c_org, const float* c_trg, int B, int T, float* mel, float* mel_post,
        float* codes, void* workspace, size_t workspace_bytes, void* stream, bool encode_only, bool bn_train = false,
        float* bn_stats = nullptr) {
    if (!h) return AVC_ERR_INVALID_ARG;
    if (h->device < 0) return fail(h, AVC_ERR_INVALID_ARG, "host-only context (device=-1) cannot run");
    if (!h->loaded) return fail(h, AVC_ERR_NOT_LOADED, "avc_load_weights has not completed");
    if (!x || !c_org || !codes || B <= 0 || T <= 0) return fail(h, AVC_ERR_INVALID_ARG, "null pointer or non-positive B/T");
    if (!encode_only && (!c_trg || !mel || !mel_post)) return fail(h, AVC_ERR_INVALID_ARG, "null c_trg/mel/mel_post");
    const avc_dims& d = h->dims;
    if (T < d.freq) return fail(h, AVC_ERR_T_TOO_SHORT, "T=%d < freq=%d: no forward content code exists (the reference raises ValueError)", T, d.freq);
    Workspace w = carve(h, (char*)workspace, B, T);
    if (!workspace || workspace_bytes < w.total || ((uintptr_t)workspace & 255))
        return fail(h, AVC_ERR_WORKSPACE, "workspace must be 256-byte aligned and >= %zu bytes (got %zu)", w.total, workspace_bytes);
    DeviceGuard guard;
    cudaError_t e = guard.enter(h->device);
    if (e != cudaSuccess) return fail(h, AVC_ERR_CUDA, "cudaSetDevice: %s", cudaGetErrorString(e));

    Runner r{h, (cudaStream_t)stream, B, T};
    r.wsp = &w;
    r.bn_train = bn_train; r.bn_stats = bn_stats;
    const bool fused_concat = !bn_train && !getenv("AVC_NO_FUSED_CONCAT");     // A/B and test aid: the unfused conv0 over the materialised concat
    const int mode = (encode_only ? 1 : 0) + (bn_train ? 3 : 0) + (fused_concat ? 0 : 8);
    h->launches = 0;
    for (auto ev : h->prof_events) cudaEventDestroy(ev);
    h->prof_events.clear(); h->prof_cats.clear();
    if (r.use_tc()) {
        r.ps = plan_set_for(h, workspace, B, T, mode, &r.building_plans);
    }

    auto enqueue = [&]() {
    const int at = act_type(h), pt = post_type(h);
    const int M = B * T;
    const int dn = d.dim_neck, E = d.dim_emb, P = d.dim_pre;
    const int nb = ceil_div(T, d.freq);

    // ---- encoder (reference encoder.py:53-97) ----
    r.mark(AVC_PHASE_LAYOUT);
    if (r.use_tc() && fused_concat && T >= 4) {
        // Fused speaker concat (reference encoder.py:65-66): cat(x, c_org expanded over t) is never built.  conv0 runs over the 80 mel
        // channels only (K = 5 x 128 instead of 5 x 384); the speaker channels' share - constant in t but for the frames whose window
        // reaches into the zero padding - is one [B, E] x [E, 5 x 512] GEMM and enters conv0's epilogue as a per-utterance bias.
        r.check(launch_pack_mel(x, c_org, w.X0, w.Cb, B, T, E, r.s), "pack mel + c_org"); h->launches++;
        r.gemm(L_ENC_CONV0_SPK, w.Cb, at, E, B, 1, ACT_NONE, w.Sspk, DT_F32, KTAPS * CONV_CH);
        r.mark(AVC_PHASE_LAYOUT);
        r.check(launch_speaker_edge(w.Sspk, h->layers[L_ENC_CONV0_MEL].b_dev, w.Espk, B, CONV_CH, r.s), "speaker edge vectors"); h->launches++;
        r.gemm(L_ENC_CONV0_MEL, w.X0, at, N_MELS, M, T, ACT_RELU, w.actA, at, CONV_CH, nullptr, nullptr, nullptr, w.Espk);
    } else {
        r.check(launch_pack_input(x, c_org, w.X0, at, B, T, E, r.s), "pack_input"); h->launches++;
        r.conv_bn(0, w.X0, at, N_MELS + E, M, T, ACT_RELU, w.actA, at, CONV_CH);
    }
    r.conv_bn(1, w.actA, at, CONV_CH, M, T, ACT_RELU, w.actB, at, CONV_CH);
    r.conv_bn(2, w.actB, at, CONV_CH, M, T, ACT_RELU, w.actA, at, CONV_CH);
    r.gemm(L_ENC_L0_IH, w.actA, at, CONV_CH, M, T, ACT_NONE, w.GxE, DT_F32, 8 * dn);
    if (r.use_tc() && !h->simt_lstm && enc_lstm_supported(dn)) {
        // register-resident BiLSTM kernels; layer 1 samples the codes and writes the decoder input itself
        r.mark(AVC_PHASE_ENC_LSTM);
        r.check(launch_enc_lstm_l0(w.GxE, h->layers[L_ENC_L0_HH].w_dev, w.Y0, B, T, r.s), "enc lstm l0"); h->launches++;
        r.gemm(L_ENC_L1_IH, w.Y0, at, 2 * dn, M, T, ACT_NONE, w.GxE, DT_F32, 8 * dn);
        r.mark(AVC_PHASE_ENC_LSTM);
        int n = 0;
        r.check(launch_enc_lstm_l1(w.GxE, h->layers[L_ENC_L1_HH].w_dev, codes, w.Dm, 2 * dn + E, encode_only ? nullptr : c_trg, E,
                                   B, T, d.freq, r.s, &n), "enc lstm l1");
        h->launches += n;
    } else {
        r.lstm(L_ENC_L0_HH, w.GxE, DT_F32, T, 1, w.Y0, at, dn, 2);
        r.gemm(L_ENC_L1_IH, w.Y0, at, 2 * dn, M, T, ACT_NONE, w.GxE, DT_F32, 8 * dn);
        r.lstm(L_ENC_L1_HH, w.GxE, DT_F32, T, 1, w.Y1, DT_F32, dn, 2);
        // code sampling + de-duplicated decoder input (encoder.py:94-95, model.py:99-124)
        r.mark(AVC_PHASE_LAYOUT);
        r.check(launch_codes(w.Y1, c_trg, codes, encode_only ? nullptr : w.Dm, at, B, T, dn, E, d.freq, r.s), "codes"); h->launches++;
    }

    if (!encode_only) {
        // ---- decoder (reference decoder.py:58-79) ----
        r.gemm(L_DEC_LSTM1_IH, w.Dm, at, 2 * dn + E, B * nb, nb, ACT_NONE, w.Gd1, DT_F32, 4 * P);
        r.lstm(L_DEC_LSTM1_HH, w.Gd1, DT_F32, nb, d.freq, w.H1, at, P, 1);
        r.conv_bn(3, w.H1, at, P, M, T, ACT_RELU, w.actA, at, P);
        r.conv_bn(4, w.actA, at, P, M, T, ACT_RELU, w.actB, at, P);
        r.conv_bn(5, w.actB, at, P, M, T, ACT_RELU, w.actA, at, P);
        const int g2t = r.gx_type(L_DEC_L2_0_HH, at, LSTM2_H);
        r.gemm(L_DEC_L2_0_IH, w.actA, at, P, M, T, ACT_NONE, w.G2, g2t, 4 * LSTM2_H);
        r.lstm(L_DEC_L2_0_HH, w.G2, g2t, T, 1, w.H2a, at, LSTM2_H, 1);
        r.gemm(L_DEC_L2_1_IH, w.H2a, at, LSTM2_H, M, T, ACT_NONE, w.G2, g2t, 4 * LSTM2_H);
        r.lstm(L_DEC_L2_1_HH, w.G2, g2t, T, 1, w.H2b, at, LSTM2_H, 1);
        // linear projection -> mel (B,80,T) fp32 + channel-last copy for the postnet (decoder.py:77, model.py:127)
        r.gemm(L_DEC_LINEAR, w.H2b, at, LSTM2_H, M, T, ACT_NONE, w.melc, pt, N_MELS, mel, nullptr);
        // ---- postnet + residual (reference postnet.py:50-63, model.py:131-132) ----
        r.conv_bn(6, w.melc, pt, N_MELS, M, T, ACT_TANH, w.actA, pt, CONV_CH);
        r.conv_bn(7, w.actA, pt, CONV_CH, M, T, ACT_TANH, w.actB, pt, CONV_CH);
        r.conv_bn(8, w.actB, pt, CONV_CH, M, T, ACT_TANH, w.actA, pt, CONV_CH);
        r.conv_bn(9, w.actA, pt, CONV_CH, M, T, ACT_TANH, w.actB, pt, CONV_CH);
        r.conv_bn(10, w.actB, pt, CONV_CH, M, T, ACT_NONE, nullptr, DT_F32, 0, mel_post, mel);
    }
    r.mark(AVC_PHASE_COUNT);  // closing marker
    };   // enqueue

    // ---- one CUDA graph per (plan set, caller pointers) ----
    // The forward is ~27 launches.  Enqueued one by one they cost the stream ~60 us of launch boundaries per step and - measured,
    // profiles/r2_e2e_scaling.md - every boundary stalls further while another stream (the host pipeline's copies) has work in
    // flight: a 21 MB device->host copy beside the step cost 0.18 ms.  Replayed as ONE graph the step is 2.28 ms instead of 2.34
    // and indifferent to concurrent copies.  So: the second call that repeats a call's pointers (steady-state loops do - a
    // pipeline's slots, or torch's allocator handing the same blocks back) captures the sequence (thread-local capture mode, on a
    // private stream), instantiates it and launches it on the caller's stream; later calls just launch it.  Not while plans are being built,
    // not under the caller's own capture, not with per-phase profiling or the step trace on.  AVC_NO_GRAPH=1 turns it off.
    PlanSet::GraphEntry* ge = nullptr;
    if (r.ps && !r.building_plans && !h->profiling && !h->no_graph) {
        cudaStreamCaptureStatus cst = cudaStreamCaptureStatusNone;
        if (cudaStreamIsCapturing(r.s, &cst) == cudaSuccess && cst == cudaStreamCaptureStatusNone) {
            const void* key[7] = {x, c_org, c_trg, mel, mel_post, codes, bn_stats};
            for (auto& g : r.ps->graphs)
                if (!memcmp(g.key, key, sizeof key)) { ge = &g; break; }
            if (!ge) {
                if (r.ps->graphs.size() >= GRAPHS_PER_SET) {
                    size_t lru = 0;
                    for (size_t i = 1; i < r.ps->graphs.size(); ++i)
                        if (r.ps->graphs[i].stamp < r.ps->graphs[lru].stamp) lru = i;
                    if (r.ps->graphs[lru].exec) cudaGraphExecDestroy(r.ps->graphs[lru].exec);
                    r.ps->graphs.erase(r.ps->graphs.begin() + lru);
                }
                r.ps->graphs.emplace_back();
                ge = &r.ps->graphs.back();
                memcpy(ge->key, key, sizeof key);
            }
            ge->stamp = ++h->plan_clock;
            ge->seen++;
        }
    }
    if (ge && ge->exec) {
        e = cudaGraphLaunch(ge->exec, r.s);
        if (e != cudaSuccess) return fail(h, AVC_ERR_CUDA, "cudaGraphLaunch: %s", cudaGetErrorString(e));
        h->launches = ge->launches;
        h->graph_replays++; r.ps->replays++;
        return AVC_OK;
    }
    // Building a graph costs a few milliseconds.  A caller whose pointers never settle (more live pointer sets than the cache
    // holds, results kept alive so that every call gets new blocks) would pay it again and again for graphs that are evicted
    // before they are reused: builds are rationed - three up front, then one more per eight replays the set's graphs have served.
    if (ge && ge->seen >= 2 && !ge->failed && r.ps->captures < 3 + r.ps->replays / 8) {
        r.ps->captures++;
        // captured on a private stream of the context (the caller's may be the legacy default stream, which cannot capture; and
        // its state is none of our business), launched on the caller's
        cudaGraph_t graph = nullptr;
        const cudaStream_t user_stream = r.s;
        e = h->capture_stream ? cudaSuccess : cudaStreamCreateWithFlags(&h->capture_stream, cudaStreamNonBlocking);
        if (e == cudaSuccess) e = cudaStreamBeginCapture(h->capture_stream, cudaStreamCaptureModeThreadLocal);
        if (e == cudaSuccess) {
            r.s = h->capture_stream;
            enqueue();
            r.s = user_stream;
            e = cudaStreamEndCapture(h->capture_stream, &graph);   // always ends the capture, also after a failed launch inside it
            if (e == cudaSuccess && r.ok() && graph) e = cudaGraphInstantiate(&ge->exec, graph, 0);
            else if (e == cudaSuccess) e = cudaErrorUnknown;
            if (graph) cudaGraphDestroy(graph);
        }
        if (e == cudaSuccess && ge->exec) {
            ge->launches = h->launches;
            e = cudaGraphLaunch(ge->exec, r.s);
            if (e != cudaSuccess) return fail(h, AVC_ERR_CUDA, "cudaGraphLaunch: %s", cudaGetErrorString(e));
            h->graph_replays++; r.ps->replays++;
            return AVC_OK;
        }
        // capture failed: nothing ran.  Forget the error, remember not to try again, and fall through to the plain sequence.
        (void)cudaGetLastError();
        ge->failed = true; ge->exec = nullptr;
        r.rc = AVC_OK; h->err.clear(); h->launches = 0;
        r.plan_cursor = r.llplan_cursor = 0; r.cur_cat = -1;
    }
    enqueue();
    if (!r.ok() && r.building_plans && r.ps) drop_plan_set(h, r.ps);   // do not keep a half-built plan set
    return r.rc;
}


// ---------------------------------------------------------------------------------------------
// d-vector speaker encoder (reference autovc/speaker_encoder/model.py:70-90): same GEMM + recurrence kernels
// ---------------------------------------------------------------------------------------------
struct SeWorkspace { char* Xp; float* Gx; char* Ya; char* Yb2; float* E; Workspace lstm; size_t total; };

SeWorkspace carve_se(const avc_ctx* h, char* base, int N, int T) {
    const size_t as = dtype_size(act_type(h));
    const size_t M = (size_t)N * T;
    SeWorkspace w{};
    size_t off = 0;
    auto take = [&](size_t bytes) { char* p = base ? base + off : nullptr; off += align256(bytes); return p; };
    w.Xp = take(M * round_up(SE_MELS, 16) * as);
    w.Gx = (float*)take(M * 4 * SE_H * 4);
    w.Ya = take(M * SE_H * as);
    w.Yb2 = take(M * SE_H * as);
    w.E = (float*)take((size_t)N * SE_EMB * 4);
    w.lstm.cstate = (float*)take((size_t)2 * N * SE_H * 4);
    w.lstm.Yb = take(ll_lstm_exchange_bytes(N, T, SE_H));
    w.total = off;
    return w;
}

int run_se(avc_ctx* h, const float* partials, int N, int T, float* embeds, void* workspace, size_t workspace_bytes, void* stream) {
    if (!h) return AVC_ERR_INVALID_ARG;
    if (h->device < 0) return fail(h, AVC_ERR_INVALID_ARG, "host-only context (device=-1) cannot run");
    if (!h->se_loaded) return fail(h, AVC_ERR_NOT_LOADED, "avc_se_load_weights has not completed");
    if (!partials || !embeds || N <= 0 || T <= 0) return fail(h, AVC_ERR_INVALID_ARG, "null pointer or non-positive N/T");
    SeWorkspace w = carve_se(h, (char*)workspace, N, T);
    if (!workspace || workspace_bytes < w.total || ((uintptr_t)workspace & 255))
        return fail(h, AVC_ERR_WORKSPACE, "workspace must be 256-byte aligned and >= %zu bytes (got %zu)", w.total, workspace_bytes);
    DeviceGuard guard;
    cudaError_t e = guard.enter(h->device);
    if (e != cudaSuccess) return fail(h, AVC_ERR_CUDA, "cudaSetDevice: %s", cudaGetErrorString(e));

    Runner r{h, (cudaStream_t)stream, N, T};
    r.wsp = &w.lstm;
    h->launches = 0;
    for (auto ev : h->prof_events) cudaEventDestroy(ev);
    h->prof_events.clear(); h->prof_cats.clear();
    if (r.use_tc()) {
        r.ps = plan_set_for(h, workspace, N, T, 2, &r.building_plans);
    }
    const int at = act_type(h);
    const int M = N * T, Kp = round_up(SE_MELS, 16);
    r.check(launch_cast_pad_rows(partials, w.Xp, at, (long long)M, SE_MELS, Kp, r.s), "cast partials"); h->launches++;
    char* in = w.Xp; int ldin = Kp;
    char* bufs[2] = {w.Ya, w.Yb2};
    for (int l = 0; l < SE_LAYERS; ++l) {
        r.gemm(L_SE_L0_IH + 2 * l, in, at, ldin, M, T, ACT_NONE, w.Gx, DT_F32, 4 * SE_H);
        r.lstm(L_SE_L0_HH + 2 * l, w.Gx, DT_F32, T, 1, bufs[l & 1], at, SE_H, 1);
        in = bufs[l & 1]; ldin = SE_H;
    }
    // hidden[-1] = last layer's h at the final frame (model.py:83-85): rows of pitch T*H starting at frame T-1
    const char* h_last = in + (size_t)(T - 1) * SE_H * dtype_size(at);
    r.gemm(L_SE_LINEAR, h_last, at, T * SE_H, N, 1, ACT_RELU, w.E, DT_F32, SE_EMB);
    r.check(launch_l2norm_rows(w.E, embeds, N, SE_EMB, r.s), "l2 normalise"); h->launches++;
    if (!r.ok() && r.building_plans && r.ps) drop_plan_set(h, r.ps);
    return r.rc;
}

}  // namespace

// ---------------------------------------------------------------------------------------------
// extern "C"
// ---------------------------------------------------------------------------------------------
// ---------------------------------------------------------------------------------------------
// allocation-free variant of avc_se_forward (no counterpart in the reference)
// ---------------------------------------------------------------------------------------------
// The scratch avc_se_forward takes from the caller lives here instead: one file-scope device array per GPU that loads the library
// (held until the process exits).  In HOST code the name below is a host-side stand-in for the symbol, not a device address:
// the address comes from cudaGetSymbolAddress, once per device, with that device current.
static __device__ __align__(256) unsigned char g_se_static_scratch[AVC_SE_STATIC_SCRATCH_BYTES];
namespace {
constexpr int MAX_DEVICES = 64;
void* g_se_static_ptr[MAX_DEVICES] = {};      // per device: device address of g_se_static_scratch (nullptr: not resolved yet)
}  // namespace

extern "C" {

int avc_abi_version(void) { return 8; }

int avc_create(avc_handle* out, const avc_dims* dims, int device, int precision) {
    if (!out || !dims) return fail(nullptr, AVC_ERR_INVALID_ARG, "null out/dims");
    *out = nullptr;
    if (precision != AVC_PREC_BF16 && precision != AVC_PREC_FP32) return fail(nullptr, AVC_ERR_INVALID_ARG, "precision must be 0 (bf16) or 1 (fp32)");
    if (dims->dim_neck <= 0 || dims->dim_emb <= 0 || dims->dim_pre <= 0 || dims->freq <= 0)
        return fail(nullptr, AVC_ERR_INVALID_ARG, "dims must be positive");
    // kernel constraints: 16-byte vector loads along K and unit-major gate tiles
    if (dims->dim_neck % 8 || dims->dim_emb % 16 || dims->dim_pre % 64)
        return fail(nullptr, AVC_ERR_UNSUPPORTED, "need dim_neck %% 8 == 0, dim_emb %% 16 == 0, dim_pre %% 64 == 0 (got %d/%d/%d)",
                    dims->dim_neck, dims->dim_emb, dims->dim_pre);
    avc_ctx* h = new (std::nothrow) avc_ctx();
    if (!h) return fail(nullptr, AVC_ERR_INVALID_ARG, "out of host memory");
    h->dims = *dims; h->device = device; h->precision = precision;
    const char* fs = getenv("AVC_FORCE_SIMT");
    h->force_simt = fs && fs[0] == '1';
    const char* sl = getenv("AVC_SIMT_LSTM");
    h->simt_lstm = sl && sl[0] == '1';
    h->gx_f32 = getenv("AVC_GX_F32") != nullptr;
    h->no_graph = getenv("AVC_NO_GRAPH") != nullptr || getenv("AVC_LSTM_TRACE") != nullptr;
    for (int i = 0; i < L_COUNT; ++i) h->layers[i].name = kLayerNames[i];
    if (device >= 0) {
        DeviceGuard guard;
        cudaError_t e = guard.enter(device);
        cudaDeviceProp prop{};
        if (e == cudaSuccess) e = cudaGetDeviceProperties(&prop, device);
        if (e != cudaSuccess) { int rc = fail(nullptr, AVC_ERR_CUDA, "cuda device %d: %s", device, cudaGetErrorString(e)); delete h; return rc; }
        if (prop.major != 10) { int rc = fail(nullptr, AVC_ERR_UNSUPPORTED, "device %d is sm_%d%d; this library is built for sm_100a (B200) only", device, prop.major, prop.minor); delete h; return rc; }
        h->num_sms = prop.multiProcessorCount;
        if (precision == AVC_PREC_BF16 && !h->force_simt) {
            const char* why = tc_init();
            if (!why) why = ll_lstm_init();
            if (why) { int rc = fail(nullptr, AVC_ERR_CUDA, "tensor-core kernel setup: %s", why); delete h; return rc; }
        }
    }
    *out = h;
    return AVC_OK;
}

int avc_load_weights(avc_handle h, const avc_weight_desc* table, int n) {
    if (!h || !table || n <= 0) return fail(h, AVC_ERR_INVALID_ARG, "null handle/table");
    for (int i = 0; i < n; ++i) {
        if (!table[i].name || !table[i].data || table[i].numel < 0) return fail(h, AVC_ERR_INVALID_ARG, "weight table entry %d is malformed", i);
        std::string name = table[i].name;
        if (name.size() > 19 && name.compare(name.size() - 19, 19, "num_batches_tracked") == 0) continue;
        h->raw[name].assign(table[i].data, table[i].data + table[i].numel);
    }
    const avc_dims& d = h->dims;
    const int dn = d.dim_neck, E = d.dim_emb, P = d.dim_pre;
    Layer* L = h->layers;
    bool ok = true;
    ok = ok && pack_conv(h, L[L_ENC_CONV0], "encoder.convolutions.0", N_MELS + E, CONV_CH);
    ok = ok && pack_conv(h, L[L_ENC_CONV1], "encoder.convolutions.1", CONV_CH, CONV_CH);
    ok = ok && pack_conv(h, L[L_ENC_CONV2], "encoder.convolutions.2", CONV_CH, CONV_CH);
    ok = ok && pack_lstm(h, L[L_ENC_L0_IH], L[L_ENC_L0_HH], "encoder.lstm", 0, CONV_CH, dn, true);
    ok = ok && pack_lstm(h, L[L_ENC_L1_IH], L[L_ENC_L1_HH], "encoder.lstm", 1, 2 * dn, dn, true);
    ok = ok && pack_lstm(h, L[L_DEC_LSTM1_IH], L[L_DEC_LSTM1_HH], "decoder.lstm1", 0, 2 * dn + E, P, false, ll_eligible(h, P));
    ok = ok && pack_conv(h, L[L_DEC_CONV0], "decoder.convolutions.0", P, P);
    ok = ok && pack_conv(h, L[L_DEC_CONV1], "decoder.convolutions.1", P, P);
    ok = ok && pack_conv(h, L[L_DEC_CONV2], "decoder.convolutions.2", P, P);
    ok = ok && pack_lstm(h, L[L_DEC_L2_0_IH], L[L_DEC_L2_0_HH], "decoder.lstm2", 0, P, LSTM2_H, false, ll_eligible(h, LSTM2_H));
    ok = ok && pack_lstm(h, L[L_DEC_L2_1_IH], L[L_DEC_L2_1_HH], "decoder.lstm2", 1, LSTM2_H, LSTM2_H, false, ll_eligible(h, LSTM2_H));
    ok = ok && pack_linear(h, L[L_DEC_LINEAR], "decoder.linear_projection.linear_layer", LSTM2_H, N_MELS);
    ok = ok && pack_conv(h, L[L_POST_CONV0], "postnet.convolutions.0", N_MELS, CONV_CH);
    ok = ok && pack_conv(h, L[L_POST_CONV1], "postnet.convolutions.1", CONV_CH, CONV_CH);
    ok = ok && pack_conv(h, L[L_POST_CONV2], "postnet.convolutions.2", CONV_CH, CONV_CH);
    ok = ok && pack_conv(h, L[L_POST_CONV3], "postnet.convolutions.3", CONV_CH, CONV_CH);
    ok = ok && pack_conv(h, L[L_POST_CONV4], "postnet.convolutions.4", CONV_CH, N_MELS);
    // the same convolutions without the fold + the BatchNorm affine parameters, for the batch-statistics forward
    static const char* kBnPrefix[N_BN] = {"encoder.convolutions.0", "encoder.convolutions.1", "encoder.convolutions.2",
                                          "decoder.convolutions.0", "decoder.convolutions.1", "decoder.convolutions.2",
                                          "postnet.convolutions.0", "postnet.convolutions.1", "postnet.convolutions.2",
                                          "postnet.convolutions.3", "postnet.convolutions.4"};
    std::vector<float> bn_gamma[N_BN], bn_beta[N_BN];
    for (int i = 0; ok && i < N_BN; ++i) {
        const Layer& F = L[kBnFolded[i]];
        ok = pack_conv(h, L[L_RAW_CONV0 + i], kBnPrefix[i], F.Cin, F.N, false);
        if (F.N > BN_MAX_CH) { ok = false; fail(h, AVC_ERR_UNSUPPORTED, "conv width %d exceeds %d", F.N, BN_MAX_CH); }
        if (ok) { bn_gamma[i] = h->raw[std::string(kBnPrefix[i]) + ".1.weight"]; bn_beta[i] = h->raw[std::string(kBnPrefix[i]) + ".1.bias"]; }
    }
    if (!ok) { h->loaded = false; return AVC_ERR_MISSING_WEIGHT; }
    split_conv0(L[L_ENC_CONV0], L[L_ENC_CONV0_MEL], L[L_ENC_CONV0_SPK], N_MELS, E);
    h->raw.clear();
    for (int i = 0; i < L_GEN_COUNT; ++i) {
        const bool post = i >= L_POST_CONV0;
        L[i].w_type = h->precision == AVC_PREC_FP32 ? DT_F32 : (post ? DT_F16 : DT_BF16);
    }
    for (int i = 0; i < N_BN; ++i) L[L_RAW_CONV0 + i].w_type = L[kBnFolded[i]].w_type;
    L[L_ENC_CONV0_MEL].w_type = L[L_ENC_CONV0_SPK].w_type = L[L_ENC_CONV0].w_type;
    if (h->device >= 0) {
        DeviceGuard guard;
        cudaError_t e = guard.enter(h->device);
        if (e != cudaSuccess) return fail(h, AVC_ERR_CUDA, "cudaSetDevice: %s", cudaGetErrorString(e));
        drop_plans(h, 0);     // cached generator plans point at the old weight buffers
        for (int i = 0; i < L_GEN_COUNT; ++i) {
            int rc = upload_layer(h, L[i]);
            if (rc != AVC_OK) return rc;
        }
        for (int i = L_ENC_CONV0_MEL; i <= L_ENC_CONV0_SPK; ++i) {
            int rc = upload_layer(h, L[i]);
            if (rc != AVC_OK) return rc;
        }
        for (int i = 0; i < N_BN; ++i) {
            int rc = upload_layer(h, L[L_RAW_CONV0 + i]);
            if (rc != AVC_OK) return rc;
            for (int k = 0; k < 2; ++k) {
                float*& dst = k ? h->bn_beta_dev[i] : h->bn_gamma_dev[i];
                const std::vector<float>& src = k ? bn_beta[i] : bn_gamma[i];
                if (dst) { cudaFree(dst); dst = nullptr; }
                e = cudaMalloc((void**)&dst, src.size() * 4);
                if (e == cudaSuccess) e = cudaMemcpy(dst, src.data(), src.size() * 4, cudaMemcpyHostToDevice);
                if (e != cudaSuccess) return fail(h, AVC_ERR_CUDA, "uploading BatchNorm parameters: %s", cudaGetErrorString(e));
            }
        }
    }
    h->loaded = true;
    return AVC_OK;
}

size_t avc_workspace_bytes(avc_handle h, int B, int T) {
    if (!h || B <= 0 || T <= 0) return 0;
    return carve(h, nullptr, B, T).total;
}

int avc_forward(avc_handle h, const float* x, const float* c_org, const float* c_trg, int B, int T, float* mel,
                float* mel_post, float* codes, void* workspace, size_t workspace_bytes, void* stream) {
    return run(h, x, c_org, c_trg, B, T, mel, mel_post, codes, workspace, workspace_bytes, stream, false);
}

int avc_encode(avc_handle h, const float* x, const float* c_org, int B, int T, float* codes, void* workspace,
               size_t workspace_bytes, void* stream) {
    return run(h, x, c_org, nullptr, B, T, nullptr, nullptr, codes, workspace, workspace_bytes, stream, true);
}

int avc_forward_bn_train(avc_handle h, const float* x, const float* c_org, const float* c_trg, int B, int T, float* mel,
                         float* mel_post, float* codes, float* bn_stats, void* workspace, size_t workspace_bytes, void* stream) {
    const bool encode_only = c_trg == nullptr;
    return run(h, x, c_org, c_trg, B, T, mel, mel_post, codes, workspace, workspace_bytes, stream, encode_only, true, bn_stats);
}

int64_t avc_last_launch_count(avc_handle h) { return h ? h->launches : 0; }

int64_t avc_plan_builds(avc_handle h) { return h ? h->plan_builds : 0; }

int64_t avc_graph_replays(avc_handle h) { return h ? h->graph_replays : 0; }

int avc_set_profiling(avc_handle h, int enable) {
    if (!h) return AVC_ERR_INVALID_ARG;
    h->profiling = enable != 0;
    return AVC_OK;
}

int avc_get_profile(avc_handle h, float* ms, int n) {
    if (!h || !ms || n < AVC_PHASE_COUNT) return fail(h, AVC_ERR_INVALID_ARG, "need room for %d phases", (int)AVC_PHASE_COUNT);
    for (int i = 0; i < n; ++i) ms[i] = 0.0f;
    if (h->prof_events.size() < 2) return fail(h, AVC_ERR_INVALID_ARG, "no profiled run recorded");
    cudaError_t e = cudaEventSynchronize(h->prof_events.back());
    if (e != cudaSuccess) return fail(h, AVC_ERR_CUDA, "cudaEventSynchronize: %s", cudaGetErrorString(e));
    for (size_t i = 0; i + 1 < h->prof_events.size(); ++i) {
        float t = 0.0f;
        cudaEventElapsedTime(&t, h->prof_events[i], h->prof_events[i + 1]);
        ms[h->prof_cats[i]] += t;
    }
    return AVC_PHASE_COUNT;
}

int64_t avc_debug_packed(avc_handle h, const char* which, float* out, int64_t cap) {
    if (!h || !which) return AVC_ERR_INVALID_ARG;
    if (!h->loaded) return fail(h, AVC_ERR_NOT_LOADED, "weights not loaded");
    std::string s = which;
    if (s.size() < 3) return fail(h, AVC_ERR_INVALID_ARG, "bad blob name '%s'", which);
    const std::string lname = s.substr(0, s.size() - 2), kind = s.substr(s.size() - 2);
    for (int i = 0; i < L_COUNT; ++i) {
        if (h->layers[i].name != lname) continue;
        const std::vector<float>& v = kind == ".w" ? h->layers[i].w : h->layers[i].b;
        if (kind != ".w" && kind != ".b") break;
        if (out) memcpy(out, v.data(), sizeof(float) * (size_t)std::min<int64_t>(cap, (int64_t)v.size()));
        return (int64_t)v.size();
    }
    return fail(h, AVC_ERR_INVALID_ARG, "unknown packed blob '%s'", which);
}

int avc_debug_gemm(avc_handle h, const char* layer, int use_tc, const void* A, int a_type, int lda, int M, int T, int act,
                   void* out, int out_type, int ldo, float* out_t, const float* res_t, void* stream) {
    if (!h || !layer || !A) return fail(h, AVC_ERR_INVALID_ARG, "null argument");
    if (!h->loaded || h->device < 0) return fail(h, AVC_ERR_NOT_LOADED, "weights not loaded on a device");
    for (int i = 0; i < L_COUNT; ++i) {
        const Layer& L = h->layers[i];
        if (L.name != layer) continue;
        GemmParams p{};
        p.A = A; p.a_type = a_type; p.lda = lda; p.M = M; p.N = L.N; p.K = L.Kp; p.T = T;
        p.taps = L.taps; p.Cin = L.Cin; p.Cp = L.Cp; p.W = L.w_dev; p.w_type = L.w_type; p.ldw = L.Kp;
        p.bias = L.b_dev; p.act = act; p.out = out; p.out_type = out_type; p.ldo = ldo; p.out_t = out_t; p.res_t = res_t;
        cudaError_t e;
        if (use_tc) {
            const char* why = tc_init();
            if (why) return fail(h, AVC_ERR_CUDA, "%s", why);
            TcGemmPlan* plan = tc_gemm_plan_create(p, h->num_sms, &why);
            if (!plan) return fail(h, AVC_ERR_UNSUPPORTED, "tensor-core GEMM plan for %s: %s", layer, why ? why : "?");
            e = launch_tc_gemm(plan, (cudaStream_t)stream, out_t, res_t);
            tc_gemm_plan_destroy(plan);   // tensor maps are passed by value at launch
        } else {
            e = launch_simt_gemm(p, (cudaStream_t)stream);
        }
        if (e != cudaSuccess) return fail(h, AVC_ERR_CUDA, "%s: %s", layer, cudaGetErrorString(e));
        return AVC_OK;
    }
    return fail(h, AVC_ERR_INVALID_ARG, "unknown layer '%s'", layer);
}

int avc_se_load_weights(avc_handle h, const avc_weight_desc* table, int n) {
    if (!h || !table || n <= 0) return fail(h, AVC_ERR_INVALID_ARG, "null handle/table");
    h->se_loaded = false;
    h->raw.clear();
    for (int i = 0; i < n; ++i) {
        if (!table[i].name || !table[i].data || table[i].numel < 0) return fail(h, AVC_ERR_INVALID_ARG, "weight table entry %d is malformed", i);
        h->raw[table[i].name].assign(table[i].data, table[i].data + table[i].numel);
    }
    Layer* L = h->layers;
    bool ok = true;
    for (int l = 0; l < SE_LAYERS; ++l)
        ok = ok && pack_lstm(h, L[L_SE_L0_IH + 2 * l], L[L_SE_L0_HH + 2 * l], "lstm", l, l == 0 ? SE_MELS : SE_H, SE_H, false, ll_eligible(h, SE_H));
    ok = ok && pack_linear(h, L[L_SE_LINEAR], "linear", SE_H, SE_EMB);
    h->raw.clear();
    if (!ok) return AVC_ERR_MISSING_WEIGHT;
    for (int i = L_SE_L0_IH; i <= L_SE_LINEAR; ++i) L[i].w_type = h->precision == AVC_PREC_FP32 ? DT_F32 : DT_BF16;
    if (h->device >= 0) {
        DeviceGuard guard;
        cudaError_t e = guard.enter(h->device);
        if (e != cudaSuccess) return fail(h, AVC_ERR_CUDA, "cudaSetDevice: %s", cudaGetErrorString(e));
        drop_plans(h, 2);     // cached speaker-encoder plans point at the old weight buffers
        for (int i = L_SE_L0_IH; i <= L_SE_LINEAR; ++i) {
            int rc = upload_layer(h, L[i]);
            if (rc != AVC_OK) return rc;
        }
    }
    h->se_loaded = true;
    return AVC_OK;
}

size_t avc_se_workspace_bytes(avc_handle h, int N, int T) {
    if (!h || N <= 0 || T <= 0) return 0;
    return carve_se(h, nullptr, N, T).total;
}

int avc_se_forward(avc_handle h, const float* partials, int N, int T, float* embeds, void* workspace, size_t workspace_bytes,
                   void* stream) {
    return run_se(h, partials, N, T, embeds, workspace, workspace_bytes, stream);
}

int avc_se_forward_static(avc_handle h, const float* partials, int N, int T, float* embeds, void* stream) {
    if (!h) return AVC_ERR_INVALID_ARG;
    if (h->device < 0 || h->device >= MAX_DEVICES) return fail(h, AVC_ERR_INVALID_ARG, "avc_se_forward_static needs a context on device 0..%d", MAX_DEVICES - 1);
    if (N <= 0 || T <= 0) return fail(h, AVC_ERR_INVALID_ARG, "non-positive N/T");
    const size_t need = carve_se(h, nullptr, N, T).total;
    if (need > (size_t)AVC_SE_STATIC_SCRATCH_BYTES)
        return fail(h, AVC_ERR_WORKSPACE, "avc_se_forward_static: N=%d partials x T=%d frames need %zu bytes of scratch, the static array holds %zu "
                    "(e.g. 8 x 160); use avc_se_forward with a caller workspace for larger batches", N, T, need, (size_t)AVC_SE_STATIC_SCRATCH_BYTES);
    if (!g_se_static_ptr[h->device]) {         // set-up, once per device, with that device current
        DeviceGuard guard;
        cudaError_t e = guard.enter(h->device);
        void* ptr = nullptr;
        if (e == cudaSuccess) e = cudaGetSymbolAddress(&ptr, g_se_static_scratch);
        if (e != cudaSuccess || !ptr) return fail(h, AVC_ERR_CUDA, "cudaGetSymbolAddress(static scratch) on device %d: %s", h->device, cudaGetErrorString(e));
        g_se_static_ptr[h->device] = ptr;
    }
    // same schedule as avc_se_forward on the static array: no cudaMalloc / cudaFree anywhere below (plans are host objects; the
    // recurrence re-arms its exchange region itself, every other buffer is fully overwritten, so stale contents are harmless)
    return run_se(h, partials, N, T, embeds, g_se_static_ptr[h->device], (size_t)AVC_SE_STATIC_SCRATCH_BYTES, stream);
}

int avc_se_reduce(avc_handle h, const float* embeds, const int32_t* offsets, int S, float* speaker_embeds, void* stream) {
    if (!h || !embeds || !offsets || !speaker_embeds || S <= 0) return fail(h, AVC_ERR_INVALID_ARG, "null pointer or S <= 0");
    if (h->device < 0) return fail(h, AVC_ERR_INVALID_ARG, "host-only context (device=-1) cannot run");
    DeviceGuard guard;
    cudaError_t e = guard.enter(h->device);
    if (e == cudaSuccess) e = launch_segment_mean_renorm(embeds, offsets, speaker_embeds, S, SE_EMB, (cudaStream_t)stream);
    if (e != cudaSuccess) return fail(h, AVC_ERR_CUDA, "segment mean: %s", cudaGetErrorString(e));
    return AVC_OK;
}

void avc_destroy(avc_handle h) {
    if (!h) return;
    if (h->device >= 0) {
        DeviceGuard guard;
        guard.enter(h->device);
        for (int i = 0; i < L_COUNT; ++i) {
            if (h->layers[i].w_dev) cudaFree(h->layers[i].w_dev);
            if (h->layers[i].b_dev) cudaFree(h->layers[i].b_dev);
        }
        for (int i = 0; i < N_BN; ++i) {
            if (h->bn_gamma_dev[i]) cudaFree(h->bn_gamma_dev[i]);
            if (h->bn_beta_dev[i]) cudaFree(h->bn_beta_dev[i]);
        }
        drop_plans(h);
        if (h->capture_stream) cudaStreamDestroy(h->capture_stream);
        for (auto ev : h->prof_events) cudaEventDestroy(ev);
    }
    delete h;
}

const char* avc_last_error(avc_handle h) { return h ? h->err.c_str() : g_create_error.c_str(); }

}  // extern "C"
