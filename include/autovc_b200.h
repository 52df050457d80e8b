/*
 * autovc_b200 - C ABI of the B200-native AutoVC generator forward pass.
 *
 * This is the drop-in boundary for ONE path of Fagprojekt-Deep-voice-conversion/AutoVC:
 *     AutoEncoder.forward(x, c_org, c_trg) -> (mel_outputs, mel_outputs_postnet, content_codes)
 *     reference: autovc/auto_encoder/model.py:68-146 (eval-mode BatchNorm, no_grad)
 * and the encoder-only pass used by the content loss
 *     AutoEncoder.forward(x, c_org, None)  -> content_codes        (model.py:98-101, model.py:201)
 *
 * The reference exposes a Python nn.Module, not an FFI; the entry points below are what a cffi /
 * ctypes / pybind binding of that module would bind (see INTEGRATION.md for the stub).  Plain
 * pointers and sizes only, no torch types; never throws across the ABI; every function returns
 * AVC_OK (0) or a negative avc_status and records a message retrievable with avc_last_error().
 *
 * Threading: one handle per device; a handle is not re-entrant (it owns recurrent scratch state
 * inside the caller's workspace).  All work is enqueued on the caller's stream, there is no
 * hidden synchronisation in avc_forward()/avc_encode().
 */
#ifndef AUTOVC_B200_H
#define AUTOVC_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct avc_ctx* avc_handle;

/* AutoEncoder ctor arguments, reference model.py:31-38 / utils/hparams.py:17-20 (32/256/512/32). */
typedef struct avc_dims {
    int32_t dim_neck; /* bottleneck width per direction            */
    int32_t dim_emb;  /* speaker-embedding width                   */
    int32_t dim_pre;  /* decoder pre-LSTM / conv width             */
    int32_t freq;     /* code sampling period in frames            */
} avc_dims;

typedef enum avc_status {
    AVC_OK = 0,
    AVC_ERR_INVALID_ARG = -1,  /* null pointer, bad dims, B<=0 ...                                  */
    AVC_ERR_T_TOO_SHORT = -2,  /* T < freq: the reference raises ValueError in torch.cat */
    AVC_ERR_CUDA = -3,         /* a CUDA runtime/driver call failed; message has the CUDA error string */
    AVC_ERR_NOT_LOADED = -4,   /* forward before a complete avc_load_weights()                        */
    AVC_ERR_WORKSPACE = -5,    /* workspace null / too small / misaligned                            */
    AVC_ERR_MISSING_WEIGHT = -6, /* a state-dict entry is absent or has the wrong element count       */
    AVC_ERR_UNSUPPORTED = -7   /* dims outside what the kernels support                              */
} avc_status;

typedef enum avc_precision {
    AVC_PREC_BF16 = 0, /* tensor-core path: bf16 operands (fp16 in the tanh-bounded postnet), fp32 accumulate */
    AVC_PREC_FP32 = 1  /* validation path: fp32 operands and accumulation on the CUDA cores                   */
} avc_precision;

/* One state-dict entry in the reference's own naming and layout (float32, contiguous, HOST memory),
 * e.g. "encoder.convolutions.0.0.conv.weight" [512,336,5], "decoder.lstm2.weight_hh_l1" [4096,1024].
 * Names are fixed by ConvNorm/LinearNorm (reference autovc/auto_encoder/net_layers.py:13,39).
 * "*.num_batches_tracked" entries are ignored.  The data is copied; the caller keeps ownership. */
typedef struct avc_weight_desc {
    const char* name;
    const float* data;
    int64_t numel;
} avc_weight_desc;

/* Create a context on CUDA device `device` (>= 0).  device == -1 creates a HOST-ONLY context that can
 * load + pack weights (for inspection with avc_debug_packed) but cannot run.  Replaces
 * AutoEncoder.__init__ (model.py:31-64). */
int avc_create(avc_handle* out, const avc_dims* dims, int device, int precision);

/* Hand over the full state dict (any order).  Folds eval-mode BatchNorm into the convolutions
 * (w' = w*g/sqrt(var+1e-5), b' = (b-mean)*g/sqrt(var+1e-5)+beta), sums bias_ih+bias_hh, re-orders LSTM
 * gate rows to unit-major (u*4+gate), re-orders conv weights to tap-major K, converts to the context's
 * precision and uploads.  Replaces AutoEncoder.load / load_state_dict (model.py:149-156). */
int avc_load_weights(avc_handle h, const avc_weight_desc* table, int n);

/* Bytes of device scratch avc_forward/avc_encode need for a (B,T) batch.  0 on error. */
size_t avc_workspace_bytes(avc_handle h, int B, int T);

/* AutoEncoder.forward (model.py:68-146).  All pointers are DEVICE pointers owned by the caller:
 *   x        [B,80,T]  float32, channel-first, contiguous
 *   c_org    [B,dim_emb], c_trg [B,dim_emb] float32
 *   mel      [B,80,T], mel_post [B,80,T] float32 outputs
 *   codes    [B, dim_neck*(floor(T/freq)+ceil(T/freq))] float32: all forward codes then all backward codes
 *   workspace: >= avc_workspace_bytes(h,B,T) bytes, 256-byte aligned
 *   stream   : cudaStream_t (as void*), may be 0 for the legacy default stream
 * Returns AVC_ERR_T_TOO_SHORT when T < freq. */
int avc_forward(avc_handle h, const float* x, const float* c_org, const float* c_trg, int B, int T,
                float* mel, float* mel_post, float* codes, void* workspace, size_t workspace_bytes,
                void* stream);

/* AutoEncoder.forward(x, c_org, None) (model.py:98-101): encoder + code sampling only. */
int avc_encode(avc_handle h, const float* x, const float* c_org, int B, int T, float* codes,
               void* workspace, size_t workspace_bytes, void* stream);

/* The same forward with BatchNorm in TRAINING mode (batch statistics over (B,T), biased variance, eps 1e-5), i.e. what
 * the reference module computes while it is in .train() - its state after construction, in AutoEncoder.learn
 * (model.py:300-321: forward at :306, the content-code re-encode at :201) and in VoiceConverter.convert, which never
 * calls .eval().  c_trg == NULL selects the encode-only mode (mel / mel_post may be NULL then).
 *   bn_stats  optional device output [11][2][512] float32: per BatchNorm layer (encoder 0-2, decoder 0-2, postnet 0-4)
 *             the batch mean and the BIASED batch variance of each channel (first N entries of each row of 512); the
 *             caller applies the running-statistics update (momentum 0.1, unbiased variance) - the library keeps no
 *             mutable model state.  Layers that did not run (encode-only: decoder / postnet) are left untouched. */
int avc_forward_bn_train(avc_handle h, const float* x, const float* c_org, const float* c_trg, int B, int T,
                         float* mel, float* mel_post, float* codes, float* bn_stats, void* workspace,
                         size_t workspace_bytes, void* stream);

/* Number of kernel launches the last avc_forward/avc_encode enqueued (for bench.py's gpu_launches). */
int64_t avc_last_launch_count(avc_handle h);

/* Number of plan sets (tensor maps + launch geometry of one (workspace, B, T, mode)) built since avc_create.  The
 * context keeps the 8 most recently used sets, so alternating shapes - length buckets of a conversion job, a B=1 call
 * between B=256 steps, the training pair's encode-only pass - do not rebuild them (tests, diagnostics). */
int64_t avc_plan_builds(avc_handle h);

/* avc_forward / avc_encode / avc_forward_bn_train replay their whole launch sequence as ONE CUDA graph once a call repeats an
 * earlier call's pointers (x, c_org, c_trg, outputs) on the same workspace / B / T: the second such call captures the
 * sequence (thread-local capture mode, on a private stream of the context) and launches the instantiated graph on the caller's
 * stream, later ones only launch it (up to 6 pointer sets per shape, least recently used first out).  Results are identical; the step is shorter and no
 * longer slowed by copies in flight on other streams.  Calls made while the caller itself is capturing, with profiling on,
 * or with AVC_NO_GRAPH=1 in the environment take the plain sequence.  Returns the number of forwards that ran as a graph
 * launch since avc_create (tests, diagnostics). */
int64_t avc_graph_replays(avc_handle h);

/* Optional per-phase device timing of the NEXT runs: CUDA events are recorded on the caller's stream at phase
 * boundaries; avc_get_profile() synchronises on the last one and returns the milliseconds spent per phase in the
 * most recent run (ms must have room for AVC_PHASE_COUNT floats).  bench.py uses this for the roofline object. */
typedef enum avc_phase {
    AVC_PHASE_LAYOUT = 0,   /* input packing, code sampling                              */
    AVC_PHASE_CONV = 1,     /* the 11 conv1d(k=5) implicit GEMMs                         */
    AVC_PHASE_PROJ = 2,     /* hoisted LSTM input projections + the final linear layer   */
    AVC_PHASE_ENC_LSTM = 3, /* encoder BiLSTM recurrences (2 layers x 2 directions)      */
    AVC_PHASE_LSTM1 = 4,    /* decoder lstm1 recurrence                                  */
    AVC_PHASE_LSTM2 = 5,    /* decoder lstm2 recurrences (2 layers)                      */
    AVC_PHASE_COUNT = 6
} avc_phase;
int avc_set_profiling(avc_handle h, int enable);
int avc_get_profile(avc_handle h, float* ms, int n);

/* Copy a packed host-side weight blob for inspection (tests).  `which`: "<layer>.w" (float32 master
 * of the packed [N,Kp] matrix) or "<layer>.b" (float32 bias); layer names are listed in DESIGN.md.
 * Returns the element count (and copies min(count, cap) floats into out if out != NULL), or a
 * negative avc_status. */
int64_t avc_debug_packed(avc_handle h, const char* which, float* out, int64_t cap);

/* Run ONE GEMM-shaped layer (conv / projection / linear, names as in avc_debug_packed) on caller-provided device
 * buffers, on the tensor cores (use_tc != 0) or on the CUDA cores, for per-layer parity tests.
 * a_type/out_type: 0 = float32, 1 = bfloat16, 2 = float16.  A is [M][lda] channel-last rows of T-frame utterances;
 * out is [M][ldo] (nullable); out_t is the channel-first float32 [M/T][N][T] output (nullable) with optional residual. */
int avc_debug_gemm(avc_handle h, const char* layer, int use_tc, const void* A, int a_type, int lda, int M, int T, int act,
                   void* out, int out_type, int ldo, float* out_t, const float* res_t, void* stream);

/* ---- d-vector speaker encoder: the producer of c_org / c_trg -------------------------------
 * Replaces SpeakerEncoder.forward (reference autovc/speaker_encoder/model.py:70-90):
 *     LSTM(40, 256, 3 layers, batch_first) -> final h of the last layer -> relu(Linear(256,256)) -> L2 normalise
 * and the partial-utterance averaging of embed_utterance (model.py:185-186).  The weights live in the same handle as
 * the generator's; state-dict names as the reference module has them: "lstm.weight_ih_l0" [1024,40],
 * "lstm.weight_hh_l0" [1024,256], "lstm.bias_ih_l0", "lstm.bias_hh_l0" [1024] (likewise l1, l2 with K = 256),
 * "linear.weight" [256,256], "linear.bias" [256].  Other entries (similarity_weight/bias) are ignored. */
int avc_se_load_weights(avc_handle h, const avc_weight_desc* table, int n);

/* Bytes of device scratch avc_se_forward needs for N partial utterances of T frames.  0 on error. */
size_t avc_se_workspace_bytes(avc_handle h, int N, int T);

/* SpeakerEncoder.forward (model.py:70-90).  Device pointers: partials [N,T,40] float32 (batch_first, the reference's
 * (batch, n_frames, n_channels)), embeds [N,256] float32 unit-norm rows.  The reference runs one utterance's partials
 * per call (model.py:141 "TODO: handle multiple wavs"); here N may span all partials of all speakers of a work list. */
int avc_se_forward(avc_handle h, const float* partials, int N, int T, float* embeds, void* workspace,
                   size_t workspace_bytes, void* stream);

/* Allocation-free variant of avc_se_forward.  NO COUNTERPART IN THE REFERENCE: added for
 * callers that cannot allocate device memory (code under CUDA graph capture, tight serving loops, harnesses that forbid
 * cudaMalloc).  Same computation and results as avc_se_forward, but the scratch is a file-scope static __device__ array of
 * AVC_SE_STATIC_SCRATCH_BYTES inside the library (every GPU that loads the library holds it until the process exits) instead
 * of a caller workspace; the call itself performs no cudaMalloc / cudaFree.  The caller still owns partials [N,T,40] and embeds
 * [N,256] (device, float32).  Limit: avc_se_workspace_bytes(h, N, T) <= AVC_SE_STATIC_SCRATCH_BYTES - one utterance's 8 partial
 * utterances of 160 frames (the reference's embed_utterance case) need 11.4 MiB; larger requests return AVC_ERR_WORKSPACE with a
 * message naming the sizes.  All calls on a device share the one array: calls that overlap on the GPU (two streams, two
 * threads, two handles on one device) race; run them one after another.  Nothing in the array has to be cleared between
 * calls (the recurrence re-arms its exchange region at every launch, everything else is overwritten).  The first call on a
 * device resolves the array's address (cudaGetSymbolAddress): make one call before starting a graph capture. */
#define AVC_SE_STATIC_SCRATCH_BYTES (12u << 20)
int avc_se_forward_static(avc_handle h, const float* partials, int N, int T, float* embeds, void* stream);

/* embed_utterance's tail (model.py:185-186) for S speakers at once: speaker s = rows [offsets[s], offsets[s+1]) of
 * embeds; speaker_embeds[s] = mean of those rows, divided by its L2 norm.  offsets: DEVICE int32 [S+1]. */
int avc_se_reduce(avc_handle h, const float* embeds, const int32_t* offsets, int S, float* speaker_embeds, void* stream);

void avc_destroy(avc_handle h);

/* ---- mel front-end: wav -> mel spectrogram on the device -------------------------------------
 * Replaces mel_spec_auto_encoder / mel_spec_speaker_encoder (reference autovc/audio/spectrogram.py:62-142, 144-217), i.e.
 * librosa.stft + librosa.filters.mel + amp_to_db + normalize_spec, for one parameter set per handle.  Independent of
 * avc_handle: a front-end owns only its window, twiddle and filterbank tables. */
typedef struct avc_mel_ctx* avc_mel_handle;

typedef struct avc_mel_config {
    int32_t sr;             /* 22050 (generator, hparams.py:7) / 16000 (speaker encoder, :53) */
    int32_t n_fft;          /* 2048 / 400 = int(sr * 25 ms) (spectrogram.py:199) */
    int32_t hop_length;     /* 275 / 160 */
    int32_t win_length;     /* 1100 / 0 = n_fft: periodic Hann, zero-padded to n_fft and centred like librosa */
    int32_t n_mels;         /* 80 / 40 */
    float fmin;             /* 40 / 0 */
    float fmax;             /* <= 0: sr / 2 */
    int32_t power;          /* 1: |X| (generator: melspectrogram(S=abs(stft))), 2: |X|^2 (speaker encoder) */
    int32_t log_normalize;  /* 1: clip((20 log10(max(1e-5, mel)) - min_level_db) / -min_level_db, 0, 1) (spectrogram.py:14-32,53-55) */
    float min_level_db;     /* -100; >= 0 selects -100 */
    int32_t pad_reflect;    /* 1: librosa <= 0.9 centre padding (reflect); 0: zeros (librosa >= 0.10) */
    int32_t frames_major;   /* 0: out[n_mels][n_frames] (generator, (80, T)); 1: out[n_frames][n_mels] (speaker encoder, .T at :204) */
} avc_mel_config;

/* device < 0: host-only handle (filterbank construction and frame arithmetic work, avc_mel_forward fails). */
int avc_mel_create(avc_mel_handle* out, const avc_mel_config* cfg, int device);

/* librosa's centred framing: 1 + n_samples / hop_length. */
int64_t avc_mel_num_frames(avc_mel_handle h, int64_t n_samples);

/* Copy the dense (n_mels, 1 + n_fft/2) float32 filterbank (= librosa.filters.mel(..., htk=False, norm="slaney")) into
 * `out` when capacity allows; returns its element count.  Host memory. */
int64_t avc_mel_filterbank(avc_mel_handle h, float* out, int64_t capacity);

/* B utterances of n_samples float32 each (DEVICE pointer, utterance b at wav + b * wav_stride) -> out (DEVICE), utterance b
 * at out + b * n_mels * n_frames in the layout `frames_major` selects.  Enqueued on `stream`. */
int avc_mel_forward(avc_mel_handle h, const float* wav, int B, int64_t n_samples, int64_t wav_stride, float* out, void* stream);

const char* avc_mel_last_error(avc_mel_handle h);   /* h may be NULL: error of the last failed avc_mel_create on this thread */
void avc_mel_destroy(avc_mel_handle h);


/* ---- WaveRNN vocoder: mel -> waveform samples -------------------------------------------------------
 * Replaces the device work of WaveRNN.generate (reference autovc/wavernn/model.py:214-329, eval mode, mode 'MOL'):
 * UpsampleNetwork.forward (:85-92 with MelResNet :33-52 and Stretch2d :55-64), the per-sample loop (:265-310: I, two GRU
 * cells, fc1-fc3) and sample_from_discretized_mix_logistic (autovc/wavernn/distribution.py:87-123).  Folding and the
 * cross-fade (:358-456) are index arithmetic / float64 numpy on the host in the reference and stay there
 * (autovc_b200/vocoder.py); the fold itself is expressed here as row windows.  Independent of avc_handle.
 * Weights: the reference module's state dict under its own names ("I.weight" [512,113], "rnn1.weight_ih_l0" [1536,512],
 * "rnn2.weight_ih_l0" [1536,544], "fc1.weight" [512,544], "fc3.weight" [30,512], "upsample.resnet.conv_in.weight"
 * [128,80,5], "upsample.resnet.layers.<i>.conv1.weight", "...batch_norm1.running_mean", "upsample.up_layers.1.weight"
 * [1,1,1,11] ...), hyper-parameters fixed to utils/hparams.py:92-107.  bf16 operands, fp32 accumulation and state. */
typedef struct avc_voc_ctx* avc_voc_handle;

int avc_voc_create(avc_voc_handle* out, int device);      /* device < 0: host-only handle (weights can be checked, nothing runs) */
int avc_voc_load_weights(avc_voc_handle h, const avc_weight_desc* table, int n);

/* UpsampleNetwork on U utterances of T mel frames: mel DEVICE [U][80][T] float32 (un-padded: the `pad` = 2 zero frames of
 * generate, model.py:252, are applied inside) -> m_up [U][T*275][80] (the up-sampled mels, transposed as :92 returns them)
 * and aux [U][T][128] at FRAME rate (the reference's aux is this repeated 275 times along time, Stretch2d :74). */
size_t avc_voc_condition_workspace_bytes(avc_voc_handle h, int U, int T);
int avc_voc_condition(avc_voc_handle h, const float* mel, int U, int T, float* m_up, float* aux, void* workspace, void* stream);

/* The sample loop on `rows` independent rows of `steps` samples.  Row i reads the conditioning of utterance row_utt[i] from
 * position row_start[i] on and zeros past the end of the utterance (pad_tensor 'after' + fold_with_overlap); the
 * unbatched generate is one row with start 0.  row_utt / row_start: DEVICE int32 [rows].  rows <= avc_voc_max_rows().
 *   uniforms : DEVICE [steps][rows][11] float32 in (0,1): 10 for the mixture choice, 1 for the logistic
 *              (distribution.py:106,117), or NULL: a Philox4x32-10 stream keyed by `seed`, step and row
 *   forced_x : NULL, or DEVICE [steps][rows]: fed back instead of the drawn sample (teacher forcing, for parity tests)
 *   samples  : DEVICE out [rows][steps] float32 in [-1, 1]
 *   logits   : NULL, or DEVICE out [steps][rows][30] (the fc3 output of every step, for parity tests) */
int avc_voc_max_rows(void);
size_t avc_voc_generate_workspace_bytes(avc_voc_handle h, int rows, int steps);
int avc_voc_generate(avc_voc_handle h, const float* m_up, const float* aux, int U, int T, const int32_t* row_utt, const int32_t* row_start,
                     int rows, int steps, const float* uniforms, uint64_t seed, const float* forced_x, float* samples, float* logits,
                     void* workspace, void* stream);
int64_t avc_voc_last_launch_count(avc_voc_handle h);
const char* avc_voc_last_error(avc_voc_handle h);   /* h may be NULL: error of the last failed avc_voc_create on this thread */
void avc_voc_destroy(avc_voc_handle h);

/* Last error message of this handle (or of the failed avc_create when h == NULL). Never NULL. */
const char* avc_last_error(avc_handle h);

/* ABI version; bumped on any signature change. */
int avc_abi_version(void);   /* 2: speaker-encoder entry points; 3: mel front-end; 4: vocoder; 5: avc_forward_bn_train; 6: avc_plan_builds (plan LRU); 7: avc_se_forward_static; 8: avc_graph_replays (graph replay of repeated calls) */

#ifdef __cplusplus
}
#endif
#endif /* AUTOVC_B200_H */
