// Mel front-end on the device: wav -> STFT -> |.|^p -> Slaney mel filterbank -> (dB, normalise), the
// producer of the generator's x[B,80,T] and of the speaker encoder's partial utterances.
//
// Replaces, for the two fixed parameter sets of the reference (autovc/utils/hparams.py:6-15, 52-58):
//   mel_spec_auto_encoder    (autovc/audio/spectrogram.py:118-127)  librosa.stft(2048, hop 275, win 1100) -> abs -> mel(80, fmin 40)
//                                                                    -> 20 log10(max(1e-5, .)) -> clip((. + 100) / 100, 0, 1)
//   mel_spec_speaker_encoder (:197-206)                              librosa.feature.melspectrogram(n_fft 400, hop 160, 40 mels, power 2), transposed
// librosa semantics restated in oracle/mel_oracle.py (centre padding, periodic Hann zero-padded to n_fft, float64 transform
// stored as complex64, float32 filterbank).
//
// One CTA per frame.  The transform runs in FLOAT64 like the reference's (numpy's rfft): with fp32 butterflies the bins 60+ dB
// below the frame's peak are noise, and the dB scale of the output makes exactly those bins visible.  n_fft = 2^p takes the
// packed real-input FFT (N/2-point complex radix-2 in shared memory + split), any other n_fft a direct DFT over the non-zero
// window span with a shared-memory twiddle table.  HBM traffic is the wav itself (each sample is read by n_fft/hop frames,
// from L2) plus n_mels floats per frame; the kernel is latency/ALU-bound, not bandwidth-bound.
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

#include <cuda_runtime.h>

#include "../../include/autovc_b200.h"

namespace {

constexpr int MEL_THREADS = 256;

struct MelArgs {
    const float* wav; long long n_samples, wav_stride;
    float* out; long long out_stride;      // per utterance
    int n_frames, n_fft, hop, win, win_off, n_bins, n_mels, log2_half;
    int power, log_normalize, reflect, frames_major;
    float min_level_db;
    const double* window;                  // [win]
    const double2* tw_half;                // pow2: exp(-2 pi i k / (n_fft/2)), k < n_fft/4;   dft: exp(-2 pi i k / n_fft), k < n_fft
    const double2* tw_full;                // pow2: exp(-2 pi i k / n_fft), k <= n_fft/2
    const int* f_start; const int* f_len; const int* f_off; const float* f_w;   // sparse filterbank rows
};

__device__ __forceinline__ double sample_at(const MelArgs& a, const float* wav, long long idx) {
    if (idx < 0) { if (!a.reflect) return 0.0; idx = -idx; }
    else if (idx >= a.n_samples) { if (!a.reflect) return 0.0; idx = 2 * (a.n_samples - 1) - idx; }
    return (double)__ldg(wav + idx);
}

// mel projection + output transform of one frame whose |X|^p sits in shared memory
__device__ __forceinline__ void mel_epilogue(const MelArgs& a, const float* mag, float* out, int frame) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int m = warp; m < a.n_mels; m += MEL_THREADS / 32) {
        const int s = a.f_start[m], n = a.f_len[m];
        const float* w = a.f_w + a.f_off[m];
        float acc = 0.0f;
        for (int i = lane; i < n; i += 32) acc = fmaf(__ldg(w + i), mag[s + i], acc);
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
        if (lane == 0) {
            float v = acc;
            if (a.log_normalize) {
                v = 20.0f * log10f(fmaxf(1e-5f, v));                                  // amp_to_db (spectrogram.py:53-55)
                v = fminf(fmaxf((v - a.min_level_db) / -a.min_level_db, 0.0f), 1.0f);  // normalize_spec (:14-32)
            }
            out[a.frames_major ? (size_t)frame * a.n_mels + m : (size_t)m * a.n_frames + frame] = v;
        }
    }
}

// n_fft = 2^p: packed real FFT.  dynamic smem: double2 z[n_fft/2] | float mag[n_fft/2 + 1]
__global__ void __launch_bounds__(MEL_THREADS) mel_fft_kernel(const MelArgs a) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    double2* z = reinterpret_cast<double2*>(smem_raw);
    const int N2 = a.n_fft >> 1;
    float* mag = reinterpret_cast<float*>(z + N2);
    const int frame = blockIdx.x, b = blockIdx.y;
    const float* wav = a.wav + (size_t)b * a.wav_stride;
    const long long base = (long long)frame * a.hop - (a.n_fft >> 1);

    // windowed frame, even samples -> re, odd -> im, stored bit-reversed for the in-place decimation-in-time transform
    for (int n = threadIdx.x; n < N2; n += MEL_THREADS) {
        double v[2];
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const int j = 2 * n + h, jw = j - a.win_off;
            v[h] = (jw >= 0 && jw < a.win) ? sample_at(a, wav, base + j) * __ldg(a.window + jw) : 0.0;
        }
        z[__brev((unsigned)n) >> (32 - a.log2_half)] = make_double2(v[0], v[1]);
    }
    __syncthreads();
    for (int s = 1; s <= a.log2_half; ++s) {
        const int half = 1 << (s - 1), tstride = N2 >> s;
        for (int bf = threadIdx.x; bf < (N2 >> 1); bf += MEL_THREADS) {
            const int pos = bf & (half - 1), i = ((bf >> (s - 1)) << s) + pos, j = i + half;
            const double2 w = __ldg(a.tw_half + pos * tstride);
            const double2 u = z[i], t = z[j];
            const double tr = t.x * w.x - t.y * w.y, ti = t.x * w.y + t.y * w.x;
            z[i] = make_double2(u.x + tr, u.y + ti);
            z[j] = make_double2(u.x - tr, u.y - ti);
        }
        __syncthreads();
    }
    // split: X[k] = E[k] + W^k O[k],  E = (Z[k] + conj Z[N2-k]) / 2,  O = -i (Z[k] - conj Z[N2-k]) / 2
    for (int k = threadIdx.x; k <= N2; k += MEL_THREADS) {
        const double2 zk = z[k & (N2 - 1)], zc = z[(N2 - k) & (N2 - 1)];
        const double er = 0.5 * (zk.x + zc.x), ei = 0.5 * (zk.y - zc.y);
        const double dr = 0.5 * (zk.x - zc.x), di = 0.5 * (zk.y + zc.y);     // (Z[k] - conj Z[N2-k]) / 2
        const double orr = di, oi = -dr;                                       // times -i
        const double2 w = __ldg(a.tw_full + k);
        const float xr = (float)(er + orr * w.x - oi * w.y), xi = (float)(ei + orr * w.y + oi * w.x);   // complex64 like librosa
        const float m = hypotf(xr, xi);
        mag[k] = a.power == 2 ? m * m : m;
    }
    __syncthreads();
    mel_epilogue(a, mag, a.out + (size_t)b * a.out_stride, frame);
}

// any n_fft: direct DFT over the non-zero window span.  dynamic smem: double xw[win] | double2 tw[n_fft] | float mag[n_bins]
__global__ void __launch_bounds__(MEL_THREADS) mel_dft_kernel(const MelArgs a) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    double2* tw = reinterpret_cast<double2*>(smem_raw);
    double* xw = reinterpret_cast<double*>(tw + a.n_fft);
    float* mag = reinterpret_cast<float*>(xw + a.win);
    const int frame = blockIdx.x, b = blockIdx.y;
    const float* wav = a.wav + (size_t)b * a.wav_stride;
    const long long base = (long long)frame * a.hop - (a.n_fft >> 1) + a.win_off;
    for (int n = threadIdx.x; n < a.win; n += MEL_THREADS) xw[n] = sample_at(a, wav, base + n) * __ldg(a.window + n);
    for (int n = threadIdx.x; n < a.n_fft; n += MEL_THREADS) tw[n] = __ldg(a.tw_half + n);
    __syncthreads();
    for (int k = threadIdx.x; k < a.n_bins; k += MEL_THREADS) {
        double re = 0.0, im = 0.0;
        int idx = (int)(((long long)k * a.win_off) % a.n_fft);
        for (int n = 0; n < a.win; ++n) {
            const double2 w = tw[idx];
            re = fma(xw[n], w.x, re);
            im = fma(xw[n], w.y, im);
            idx += k;
            if (idx >= a.n_fft) idx -= a.n_fft;
        }
        const float m = hypotf((float)re, (float)im);
        mag[k] = a.power == 2 ? m * m : m;
    }
    __syncthreads();
    mel_epilogue(a, mag, a.out + (size_t)b * a.out_stride, frame);
}

// ---- host side: Slaney filterbank exactly as librosa.filters.mel builds it (float64 ramps, float32 weights) ----
double hz_to_mel(double f) {
    const double f_sp = 200.0 / 3, min_log_hz = 1000.0, min_log_mel = min_log_hz / f_sp, logstep = std::log(6.4) / 27.0;
    return f >= min_log_hz ? min_log_mel + std::log(f / min_log_hz) / logstep : f / f_sp;
}
double mel_to_hz(double m) {
    const double f_sp = 200.0 / 3, min_log_hz = 1000.0, min_log_mel = min_log_hz / f_sp, logstep = std::log(6.4) / 27.0;
    return m >= min_log_mel ? min_log_hz * std::exp(logstep * (m - min_log_mel)) : f_sp * m;
}

}  // namespace

struct avc_mel_ctx {
    avc_mel_config cfg{};
    int device = 0;
    bool pow2 = false;
    int n_bins = 0, log2_half = 0, win_off = 0;
    size_t smem = 0;
    double* d_window = nullptr; double2* d_tw_half = nullptr; double2* d_tw_full = nullptr;
    int* d_start = nullptr; int* d_len = nullptr; int* d_off = nullptr; float* d_w = nullptr;
    std::vector<float> dense;      // host copy of the filterbank, [n_mels][n_bins] (avc_mel_filterbank)
    std::string err;
};

namespace {
thread_local std::string g_mel_err;
int mel_fail(avc_mel_ctx* h, int code, const char* fmt, ...) {
    char buf[512];
    va_list ap; va_start(ap, fmt); vsnprintf(buf, sizeof buf, fmt, ap); va_end(ap);
    if (h) h->err = buf; else g_mel_err = buf;
    return code;
}
template <class T> cudaError_t upload(T** dst, const std::vector<T>& v) {
    cudaError_t e = cudaMalloc(dst, v.size() * sizeof(T) + 16);
    if (e != cudaSuccess) return e;
    return cudaMemcpy(*dst, v.data(), v.size() * sizeof(T), cudaMemcpyHostToDevice);
}
}  // namespace

extern "C" {

int avc_mel_create(avc_mel_handle* out, const avc_mel_config* cfg, int device) {
    if (!out || !cfg) return mel_fail(nullptr, AVC_ERR_INVALID_ARG, "null out/cfg");
    const avc_mel_config& c = *cfg;
    if (c.sr <= 0 || c.n_fft < 16 || c.n_fft > 8192 || c.hop_length <= 0 || c.n_mels <= 0 || c.n_mels > 1024 ||
        (c.power != 1 && c.power != 2) || c.win_length > c.n_fft || c.win_length < 0)
        return mel_fail(nullptr, AVC_ERR_INVALID_ARG, "bad mel configuration");
    avc_mel_ctx* h = new avc_mel_ctx();
    h->cfg = c;
    if (h->cfg.win_length == 0) h->cfg.win_length = c.n_fft;
    if (h->cfg.fmax <= 0.0f) h->cfg.fmax = 0.5f * c.sr;
    if (h->cfg.min_level_db >= 0.0f) h->cfg.min_level_db = -100.0f;
    h->device = device;
    const int n_fft = c.n_fft, win = h->cfg.win_length;
    h->n_bins = 1 + n_fft / 2;
    h->win_off = (n_fft - win) / 2;
    h->pow2 = (n_fft & (n_fft - 1)) == 0;
    for (int v = n_fft / 2; v > 1; v >>= 1) ++h->log2_half;

    // filterbank (librosa.filters.mel, htk=False, norm="slaney")
    const int n_mels = c.n_mels, nb = h->n_bins;
    std::vector<double> mel_f(n_mels + 2);
    const double m0 = hz_to_mel(c.fmin), m1 = hz_to_mel(h->cfg.fmax);
    for (int i = 0; i < n_mels + 2; ++i) {
        // numpy.linspace: start + i * step, with the last point set to stop exactly
        const double step = (m1 - m0) / (n_mels + 1);
        mel_f[i] = mel_to_hz(i == n_mels + 1 ? m1 : m0 + i * step);
    }
    h->dense.assign((size_t)n_mels * nb, 0.0f);
    std::vector<int> start(n_mels), len(n_mels), off(n_mels);
    std::vector<float> flat;
    for (int i = 0; i < n_mels; ++i) {
        const double fd0 = mel_f[i + 1] - mel_f[i], fd1 = mel_f[i + 2] - mel_f[i + 1];
        const double enorm = 2.0 / (mel_f[i + 2] - mel_f[i]);
        int first = -1, last = -1;
        for (int k = 0; k < nb; ++k) {
            const double fk = k == nb - 1 ? 0.5 * c.sr : k * ((0.5 * c.sr) / (nb - 1));
            const double lower = -(mel_f[i] - fk) / fd0, upper = (mel_f[i + 2] - fk) / fd1;
            const float w32 = (float)std::fmax(0.0, std::fmin(lower, upper));
            const float w = (float)((double)w32 * enorm);
            h->dense[(size_t)i * nb + k] = w;
            if (w != 0.0f) { if (first < 0) first = k; last = k; }
        }
        start[i] = first < 0 ? 0 : first;
        len[i] = first < 0 ? 0 : last - first + 1;
        off[i] = (int)flat.size();
        for (int k = 0; k < len[i]; ++k) flat.push_back(h->dense[(size_t)i * nb + start[i] + k]);
    }
    if (flat.empty()) flat.push_back(0.0f);

    if (device >= 0) {
        std::vector<double> window(win);
        const double pi = 3.14159265358979323846;
        for (int n = 0; n < win; ++n) window[n] = 0.5 - 0.5 * std::cos(2.0 * pi * n / win);     // periodic Hann (fftbins=True)
        std::vector<double2> twh, twf;
        if (h->pow2) {
            const int N2 = n_fft / 2;
            twh.resize(N2 / 2 > 0 ? N2 / 2 : 1);
            for (int k = 0; k < N2 / 2; ++k) twh[k] = make_double2(std::cos(2.0 * pi * k / N2), -std::sin(2.0 * pi * k / N2));
            twf.resize(N2 + 1);
            for (int k = 0; k <= N2; ++k) twf[k] = make_double2(std::cos(2.0 * pi * k / n_fft), -std::sin(2.0 * pi * k / n_fft));
            h->smem = (size_t)N2 * sizeof(double2) + (size_t)(N2 + 1) * sizeof(float);
        } else {
            twh.resize(n_fft);
            for (int k = 0; k < n_fft; ++k) twh[k] = make_double2(std::cos(2.0 * pi * k / n_fft), -std::sin(2.0 * pi * k / n_fft));
            twf.resize(1, make_double2(1.0, 0.0));
            h->smem = (size_t)n_fft * sizeof(double2) + (size_t)win * sizeof(double) + (size_t)nb * sizeof(float);
        }
        cudaError_t e = cudaSetDevice(device);
        if (e == cudaSuccess) e = upload(&h->d_window, window);
        if (e == cudaSuccess) e = upload(&h->d_tw_half, twh);
        if (e == cudaSuccess) e = upload(&h->d_tw_full, twf);
        if (e == cudaSuccess) e = upload(&h->d_start, start);
        if (e == cudaSuccess) e = upload(&h->d_len, len);
        if (e == cudaSuccess) e = upload(&h->d_off, off);
        if (e == cudaSuccess) e = upload(&h->d_w, flat);
        if (e == cudaSuccess && h->smem > 48 * 1024)
            e = cudaFuncSetAttribute(h->pow2 ? (const void*)mel_fft_kernel : (const void*)mel_dft_kernel,
                                     cudaFuncAttributeMaxDynamicSharedMemorySize, (int)h->smem);
        if (e != cudaSuccess) {
            const int rc = mel_fail(nullptr, AVC_ERR_CUDA, "mel front-end setup: %s", cudaGetErrorString(e));
            avc_mel_destroy(h);
            return rc;
        }
    }
    *out = h;
    return AVC_OK;
}

int64_t avc_mel_num_frames(avc_mel_handle h, int64_t n_samples) {
    if (!h || n_samples < 0) return -1;
    return 1 + n_samples / h->cfg.hop_length;
}

int64_t avc_mel_filterbank(avc_mel_handle h, float* out, int64_t capacity) {
    if (!h) return -1;
    const int64_t n = (int64_t)h->dense.size();
    if (out && capacity >= n) memcpy(out, h->dense.data(), (size_t)n * sizeof(float));
    return n;
}

int avc_mel_forward(avc_mel_handle h, const float* wav, int B, int64_t n_samples, int64_t wav_stride, float* out, void* stream) {
    if (!h) return AVC_ERR_INVALID_ARG;
    if (h->device < 0) return mel_fail(h, AVC_ERR_INVALID_ARG, "host-only context (device=-1) cannot run");
    if (!wav || !out || B <= 0 || n_samples <= 0 || wav_stride < n_samples) return mel_fail(h, AVC_ERR_INVALID_ARG, "null pointer or bad sizes");
    if (h->cfg.pad_reflect && n_samples <= h->cfg.n_fft / 2)
        return mel_fail(h, AVC_ERR_INVALID_ARG, "reflect padding needs more than n_fft/2 = %d samples (got %lld)", h->cfg.n_fft / 2, (long long)n_samples);
    const int64_t n_frames = 1 + n_samples / h->cfg.hop_length;
    if (n_frames > 0x7fffffff || B > 65535) return mel_fail(h, AVC_ERR_INVALID_ARG, "too many frames / utterances for one launch");
    cudaError_t e = cudaSetDevice(h->device);
    if (e != cudaSuccess) return mel_fail(h, AVC_ERR_CUDA, "cudaSetDevice: %s", cudaGetErrorString(e));
    MelArgs a{};
    a.wav = wav; a.n_samples = n_samples; a.wav_stride = wav_stride;
    a.out = out; a.out_stride = n_frames * h->cfg.n_mels;
    a.n_frames = (int)n_frames; a.n_fft = h->cfg.n_fft; a.hop = h->cfg.hop_length; a.win = h->cfg.win_length; a.win_off = h->win_off;
    a.n_bins = h->n_bins; a.n_mels = h->cfg.n_mels; a.log2_half = h->log2_half;
    a.power = h->cfg.power; a.log_normalize = h->cfg.log_normalize; a.reflect = h->cfg.pad_reflect; a.frames_major = h->cfg.frames_major;
    a.min_level_db = h->cfg.min_level_db;
    a.window = h->d_window; a.tw_half = h->d_tw_half; a.tw_full = h->d_tw_full;
    a.f_start = h->d_start; a.f_len = h->d_len; a.f_off = h->d_off; a.f_w = h->d_w;
    const dim3 grid((unsigned)n_frames, (unsigned)B);
    if (h->pow2) mel_fft_kernel<<<grid, MEL_THREADS, h->smem, (cudaStream_t)stream>>>(a);
    else mel_dft_kernel<<<grid, MEL_THREADS, h->smem, (cudaStream_t)stream>>>(a);
    e = cudaGetLastError();
    if (e != cudaSuccess) return mel_fail(h, AVC_ERR_CUDA, "mel kernel launch: %s", cudaGetErrorString(e));
    return AVC_OK;
}

const char* avc_mel_last_error(avc_mel_handle h) { return h ? h->err.c_str() : g_mel_err.c_str(); }

void avc_mel_destroy(avc_mel_handle h) {
    if (!h) return;
    if (h->device >= 0) {
        cudaSetDevice(h->device);
        cudaFree(h->d_window); cudaFree(h->d_tw_half); cudaFree(h->d_tw_full);
        cudaFree(h->d_start); cudaFree(h->d_len); cudaFree(h->d_off); cudaFree(h->d_w);
    }
    delete h;
}

}  // extern "C"
