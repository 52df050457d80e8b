"""CPU oracle of the WaveRNN vocoder's `generate` - TEST INFRASTRUCTURE, never on the product path.

Restates in numpy (float32 arithmetic, as the reference's CPU tensors) what `autovc/wavernn/model.py` computes:

    MelResNet / ResBlock      (:16-52)    conv_in k=5 -> BN -> relu -> 10 x (1x1 -> BN -> relu -> 1x1 -> BN -> +res) -> conv_out
    Stretch2d / UpsampleNetwork (:55-92)  repeat-stretch + (1, 2s+1) convolutions for s in (5, 5, 11), cropped by pad*275
    WaveRNN.generate          (:214-329)  pad, upsample, fold, the per-sample loop (I, 2 GRU cells, fc1-fc3, sampling),
                                           cross-fade unfold, fade-out
    fold_with_overlap / xfade_and_unfold (:358-456)
    sample_from_discretized_mix_logistic (`autovc/wavernn/distribution.py:87-123`)

BatchNorm runs on its running statistics: `generate` calls `self.eval()` (:239).

The reference draws its random numbers from torch's global generator; here the stream is an explicit input
(`uniforms[t, row, 0:10]` for the mixture choice, `[t, row, 10]` for the logistic, both U(1e-5, 1-1e-5)), and the
golden vectors (oracle/make_golden_wavernn.py) are produced by the unmodified reference module with `Tensor.uniform_`
patched to read the same stream - that pins this file against the reference (tests/test_wavernn_oracle.py).

`forced_x` (n_steps, rows) replaces the fed-back sample (teacher forcing): the loop is chaotic in its own samples (a
flipped mixture choice changes everything after it), so element-wise parity of a bf16 device path is checked on the
logits / samples of a forced run plus the prefix of a free run.
"""
from __future__ import annotations

import numpy as np

HOP, PAD, AUX = 275, 2, 32
FACTORS = (5, 5, 11)
LOG_SCALE_MIN = float(np.log(1e-14))


def _np(sd, k):
    v = sd[k]
    return (v.detach().cpu().numpy() if hasattr(v, "detach") else np.asarray(v)).astype(np.float32)


def _bn(sd, p, x, eps=1e-5):                    # x (B, C, T), eval mode
    g, b, m, v = (_np(sd, p + s) for s in (".weight", ".bias", ".running_mean", ".running_var"))
    return (x - m[None, :, None]) / np.sqrt(v[None, :, None] + np.float32(eps)) * g[None, :, None] + b[None, :, None]


def _conv1x1(w, x):                              # w (Co, Ci, 1), x (B, Ci, T)
    return np.einsum("oc,bct->bot", w[:, :, 0], x, optimize=True).astype(np.float32)


def mel_resnet(sd, x):                           # model.py:33-52; x (B, 80, T+4) -> (B, 128, T)
    w = _np(sd, "upsample.resnet.conv_in.weight")
    k = w.shape[2]
    T = x.shape[2] - k + 1
    y = np.zeros((x.shape[0], w.shape[0], T), dtype=np.float32)
    for j in range(k):
        y += np.einsum("oc,bct->bot", w[:, :, j], x[:, :, j:j + T], optimize=True)
    y = np.maximum(_bn(sd, "upsample.resnet.batch_norm", y), 0)
    i = 0
    while f"upsample.resnet.layers.{i}.conv1.weight" in sd:
        p = f"upsample.resnet.layers.{i}"
        r = y
        y = np.maximum(_bn(sd, p + ".batch_norm1", _conv1x1(_np(sd, p + ".conv1.weight"), y)), 0)
        y = _bn(sd, p + ".batch_norm2", _conv1x1(_np(sd, p + ".conv2.weight"), y)) + r
        i += 1
    return _conv1x1(_np(sd, "upsample.resnet.conv_out.weight"), y) + _np(sd, "upsample.resnet.conv_out.bias")[None, :, None]


def upsample(sd, m):
    """UpsampleNetwork.forward (model.py:85-92): m (B, 80, T+4) -> mels (B, T*275, 80), aux (B, T*275, 128)."""
    m = np.asarray(m, dtype=np.float32)
    aux = np.repeat(mel_resnet(sd, m), HOP, axis=2)                       # Stretch2d(total_scale, 1)
    for i, s in enumerate(FACTORS):
        w = _np(sd, f"upsample.up_layers.{2 * i + 1}.weight").reshape(-1)
        m = np.repeat(m, s, axis=2)                                         # Stretch2d(s, 1)
        mp = np.pad(m, ((0, 0), (0, 0), (s, s)))
        L = m.shape[2]
        out = np.zeros_like(m)
        for j in range(2 * s + 1):                                          # cross-correlation, zero padding s
            out += w[j] * mp[:, :, j:j + L]
        m = out
    indent = PAD * HOP
    m = m[:, :, indent:-indent]
    return np.ascontiguousarray(m.transpose(0, 2, 1)), np.ascontiguousarray(aux.transpose(0, 2, 1))


def pad_after(x, pad):                           # pad_tensor(side='after'), model.py:344-356
    return np.concatenate([x, np.zeros((x.shape[0], pad, x.shape[2]), dtype=x.dtype)], axis=1)


def fold_params(total_len, target, overlap):     # model.py:381-391 -> (num_folds, padding)
    num_folds = (total_len - overlap) // (target + overlap)
    extended_len = num_folds * (overlap + target) + overlap
    remaining = total_len - extended_len
    padding = 0
    if remaining != 0:
        num_folds += 1
        padding = target + 2 * overlap - remaining
    return num_folds, padding


def fold_with_overlap(x, target, overlap):       # model.py:358-401; x (1, L, F)
    num_folds, padding = fold_params(x.shape[1], target, overlap)
    if padding:
        x = pad_after(x, padding)
    folded = np.zeros((num_folds, target + 2 * overlap, x.shape[2]), dtype=x.dtype)
    for i in range(num_folds):
        start = i * (target + overlap)
        folded[i] = x[0, start:start + target + 2 * overlap]
    return folded


def xfade_and_unfold(y, target, overlap):        # model.py:403-456; y (num_folds, target + 2*overlap) float64, modified in place
    num_folds, length = y.shape
    target = length - 2 * overlap
    total_len = num_folds * (target + overlap) + overlap
    silence_len = overlap // 2
    fade_len = overlap - silence_len
    t = np.linspace(-1, 1, fade_len, dtype=np.float64)
    fade_in = np.concatenate([np.zeros(silence_len), np.sqrt(0.5 * (1 + t))])
    fade_out = np.concatenate([np.ones(silence_len), np.sqrt(0.5 * (1 - t))])
    y[:, :overlap] *= fade_in
    y[:, -overlap:] *= fade_out
    unfolded = np.zeros(total_len, dtype=np.float64)
    for i in range(num_folds):
        start = i * (target + overlap)
        unfolded[start:start + target + 2 * overlap] += y[i]
    return unfolded


def _sigmoid(x):
    return (1.0 / (1.0 + np.exp(-x))).astype(np.float32)


def gru_cell(x, h, w_ih, w_hh, b_ih, b_hh):      # torch.nn.GRUCell
    gi = x @ w_ih.T + b_ih
    gh = h @ w_hh.T + b_hh
    H = h.shape[1]
    r = _sigmoid(gi[:, :H] + gh[:, :H])
    z = _sigmoid(gi[:, H:2 * H] + gh[:, H:2 * H])
    n = np.tanh(gi[:, 2 * H:] + r * gh[:, 2 * H:])
    return ((1 - z) * n + z * h).astype(np.float32)


def sample_mol(logits, u):
    """distribution.py:87-123 with the random stream given: logits (B, 30), u (B, 11) -> (B,) in [-1, 1]."""
    nr = logits.shape[1] // 3
    temp = logits[:, :nr] - np.log(-np.log(u[:, :nr]))
    k = np.argmax(temp, axis=1)
    rows = np.arange(logits.shape[0])
    means = logits[rows, nr + k]
    log_scales = np.maximum(logits[rows, 2 * nr + k], np.float32(LOG_SCALE_MIN))
    uu = u[:, nr]
    x = means + np.exp(log_scales) * (np.log(uu) - np.log(np.float32(1.0) - uu))
    return np.clip(x, -1.0, 1.0).astype(np.float32)


def generate_rows(sd, mels, aux, uniforms, forced_x=None, return_logits=False):
    """The per-sample loop of `generate` (model.py:265-310) on already up-sampled / folded conditioning:
    mels (B, L, 80), aux (B, L, 128), uniforms (L, B, 11) -> samples (B, L) [, logits (L, B, 30)]."""
    W = {k: _np(sd, k) for k in ("I.weight", "I.bias", "fc1.weight", "fc1.bias", "fc2.weight", "fc2.bias", "fc3.weight", "fc3.bias")}
    g1 = [_np(sd, "rnn1." + k) for k in ("weight_ih_l0", "weight_hh_l0", "bias_ih_l0", "bias_hh_l0")]
    g2 = [_np(sd, "rnn2." + k) for k in ("weight_ih_l0", "weight_hh_l0", "bias_ih_l0", "bias_hh_l0")]
    B, L, _ = mels.shape
    R = W["I.weight"].shape[0]
    h1 = np.zeros((B, R), dtype=np.float32)
    h2 = np.zeros((B, R), dtype=np.float32)
    x = np.zeros((B, 1), dtype=np.float32)
    out = np.zeros((B, L), dtype=np.float32)
    logits_all = np.zeros((L, B, W["fc3.weight"].shape[0]), dtype=np.float32) if return_logits else None
    a = [aux[:, :, AUX * i:AUX * (i + 1)] for i in range(4)]
    for i in range(L):
        xi = np.concatenate([x, mels[:, i], a[0][:, i]], axis=1) @ W["I.weight"].T + W["I.bias"]
        h1 = gru_cell(xi, h1, *g1)
        xi = xi + h1
        h2 = gru_cell(np.concatenate([xi, a[1][:, i]], axis=1), h2, *g2)
        xi = xi + h2
        y = np.maximum(np.concatenate([xi, a[2][:, i]], axis=1) @ W["fc1.weight"].T + W["fc1.bias"], 0)
        y = np.maximum(np.concatenate([y, a[3][:, i]], axis=1) @ W["fc2.weight"].T + W["fc2.bias"], 0)
        logits = (y @ W["fc3.weight"].T + W["fc3.bias"]).astype(np.float32)
        s = sample_mol(logits, uniforms[i])
        out[:, i] = s
        if return_logits:
            logits_all[i] = logits
        x = (s if forced_x is None else forced_x[i]).reshape(B, 1).astype(np.float32)
    return (out, logits_all) if return_logits else out


def condition(sd, mel, batched=True, target=11000, overlap=550):
    """generate (model.py:248-257) up to the loop: mel (1, 80, T) -> mels (B, L, 80), aux (B, L, 128), wave_len."""
    mel = np.asarray(mel, dtype=np.float32)
    wave_len = (mel.shape[-1] - 1) * HOP
    padded = np.pad(mel, ((0, 0), (0, 0), (PAD, PAD)))
    mels, aux = upsample(sd, padded)
    if batched:
        mels = fold_with_overlap(mels, target, overlap)
        aux = fold_with_overlap(aux, target, overlap)
    return mels, aux, wave_len


def finish(samples, wave_len, batched=True, target=11000, overlap=550):
    """generate (model.py:312-329) after the loop: samples (B, L) float32 -> waveform float32 (wave_len,)."""
    output = samples.astype(np.float64)
    output = xfade_and_unfold(output, target, overlap) if batched else output[0]
    fade_out = np.linspace(1, 0, 20 * HOP)
    output = output[:wave_len]
    output[-20 * HOP:] *= fade_out
    return output.astype(np.float32)


def generate(sd, mel, uniforms, batched=True, target=11000, overlap=550):
    mels, aux, wave_len = condition(sd, mel, batched, target, overlap)
    samples = generate_rows(sd, mels, aux, uniforms)
    return finish(samples, wave_len, batched, target, overlap)
