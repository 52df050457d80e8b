"""CPU oracle for the AutoVC generator forward pass.  TEST INFRASTRUCTURE ONLY.

Only `tests/`, `__graft_entry__.smoke()` and `bench.py`'s `cpu_baseline` / `--impl reference`
legs may import this module; the product (`autovc_b200/`) never does and has no CPU fallback.

This is a *restatement* of the reference algorithm with every operator spelled out as explicit
tensor arithmetic (matmul, add, sigmoid, tanh) on CPU tensors - no `nn.Conv1d`, `nn.LSTM`,
`nn.BatchNorm1d` or `nn.Linear` - so that the kernels have an unambiguous specification:

* conv1d k=5, stride 1, zero padding 2          reference `autovc/auto_encoder/net_layers.py:39-49`
* BatchNorm1d in eval mode, eps = 1e-5          reference `encoder.py:46`, `decoder.py:48`, `postnet.py:27,38,47`
  (and in training mode on batch statistics, `batchnorm_train`: the module as `learn` / `convert` run it)
* LSTM, gate order i,f,g,o, b_ih + b_hh         reference `encoder.py:51,75`, `decoder.py:37,53,65,74`
* speaker concat, code sampling                 reference `encoder.py:65-66,94-95`
* code up-sampling, decoder, postnet, residual  reference `model.py:92-146`, `decoder.py:58-79`, `postnet.py:50-63`
* chunk overlap-average (`batch_forward`)       reference `model.py:395-437`

Parity pinning: the reference ships no tests or golden vectors for this path, so
the oracle is pinned against the reference module itself, imported in the build container by
`oracle/make_golden.py`; the resulting known-answer vectors live in `tests/golden/*.npz` and
`tests/test_oracle.py` checks this file against them (max-abs <= 2e-5 in fp32).

`quant` lets a test emulate the product's bf16 operand rounding on the CPU (to predict the bf16
error budget without a GPU); the default is exact fp32 (or fp64 with `dtype=torch.float64`).
"""
from __future__ import annotations

from typing import Callable, Dict, List, Optional, Tuple

import torch

BN_EPS = 1e-5  # nn.BatchNorm1d default; the reference constructs BatchNorm1d(512) with defaults

Tensor = torch.Tensor
Quant = Optional[Callable[[Tensor, str], Tensor]]


def bf16_round(t: Tensor, tag: str = "") -> Tensor:
    """Round-to-nearest-even to bfloat16 and back (operand rounding of the tensor-core path).
    `tag` names the operand ("<layer>:x" activation, "<layer>:w" weight, "<layer>:h" recurrent
    state) so a test can quantise layers selectively."""
    return t.to(torch.bfloat16).to(t.dtype)


def _q(t: Tensor, quant: Quant, tag: str = "") -> Tensor:
    return t if quant is None else quant(t, tag)


# --------------------------------------------------------------------------------------------
# operators
# --------------------------------------------------------------------------------------------
def conv1d_k5(x: Tensor, w: Tensor, b: Tensor, quant: Quant = None, tag: str = "conv") -> Tensor:
    """x (B,Cin,T), w (Cout,Cin,5), b (Cout) -> (B,Cout,T); zero padding 2 each side.

    out[b,co,t] = b[co] + sum_k sum_ci w[co,ci,k] * x[b,ci,t+k-2]   (net_layers.py:39-49)."""
    B, Cin, T = x.shape
    Cout, _, K = w.shape
    pad = (K - 1) // 2
    xp = torch.zeros(B, Cin, T + 2 * pad, dtype=x.dtype)
    xp[:, :, pad:pad + T] = _q(x, quant, tag + ":x")
    wq = _q(w, quant, tag + ":w")
    out = b.view(1, Cout, 1).expand(B, Cout, T).clone()
    for k in range(K):
        # (Cout,Cin) @ (B,Cin,T) -> (B,Cout,T)
        out += torch.matmul(wq[:, :, k], xp[:, :, k:k + T])
    return out


def batchnorm_eval(x: Tensor, gamma: Tensor, beta: Tensor, mean: Tensor, var: Tensor) -> Tensor:
    """y = (x - mean) / sqrt(var + eps) * gamma + beta over the channel dim of (B,C,T)."""
    inv = gamma / torch.sqrt(var + BN_EPS)
    return (x - mean.view(1, -1, 1)) * inv.view(1, -1, 1) + beta.view(1, -1, 1)


def batchnorm_train(x: Tensor, gamma: Tensor, beta: Tensor) -> Tuple[Tensor, Tensor, Tensor]:
    """nn.BatchNorm1d in training mode on (B,C,T): per-channel mean and BIASED variance over (B,T), eps 1e-5.
    Returns (y, batch mean, biased batch variance).  reference `encoder.py:46,70` with the module in `.train()`
    (its state in `AutoEncoder.learn`, `model.py:306,201`, and in the un-eval()'d `convert`)."""
    mean = x.mean(dim=(0, 2))
    var = ((x - mean.view(1, -1, 1)) ** 2).mean(dim=(0, 2))
    y = (x - mean.view(1, -1, 1)) / torch.sqrt(var.view(1, -1, 1) + BN_EPS) * gamma.view(1, -1, 1) + beta.view(1, -1, 1)
    return y, mean, var


def running_stats_update(running_mean: Tensor, running_var: Tensor, mean: Tensor, var_biased: Tensor, n: int,
                         momentum: float = 0.1) -> Tuple[Tensor, Tensor]:
    """The buffer update nn.BatchNorm1d performs after a training-mode forward over n = B*T values per channel:
    running = (1-m)*running + m*stat, with the UNBIASED variance n/(n-1)*var."""
    return ((1 - momentum) * running_mean + momentum * mean,
            (1 - momentum) * running_var + momentum * var_biased * (n / (n - 1)))


def fold_batchnorm(w: Tensor, b: Tensor, gamma: Tensor, beta: Tensor, mean: Tensor, var: Tensor):
    """Fold eval-mode BN into the conv: w' = w*s, b' = (b-mean)*s + beta, s = gamma/sqrt(var+eps)."""
    s = gamma / torch.sqrt(var + BN_EPS)
    return w * s.view(-1, 1, 1), (b - mean) * s + beta


def lstm_direction(x: Tensor, w_ih: Tensor, w_hh: Tensor, b_ih: Tensor, b_hh: Tensor,
                   reverse: bool = False, quant: Quant = None, tag: str = "lstm") -> Tuple[Tensor, Tensor]:
    """One LSTM direction over x (B,T,I) -> (outputs (B,T,H), final h (B,H)).

    PyTorch equations (gate order i,f,g,o):  g = x W_ih^T + b_ih + h W_hh^T + b_hh;
    c' = sig(f) c + sig(i) tanh(g);  h' = sig(o) tanh(c').  The reverse direction walks
    t = T-1..0 and stores its output at its own t."""
    B, T, _ = x.shape
    H = w_hh.shape[1]
    gx = torch.matmul(_q(x, quant, tag + ".ih:x"), _q(w_ih, quant, tag + ".ih:w").t()) + (b_ih + b_hh)  # hoisted input projection
    whh_t = _q(w_hh, quant, tag + ".hh:w").t().contiguous()
    h = torch.zeros(B, H, dtype=x.dtype)
    c = torch.zeros(B, H, dtype=x.dtype)
    out = torch.empty(B, T, H, dtype=x.dtype)
    steps = range(T - 1, -1, -1) if reverse else range(T)
    for t in steps:
        g = gx[:, t, :] + torch.matmul(_q(h, quant, tag + ".hh:h"), whh_t)
        i, f, gg, o = g[:, :H], g[:, H:2 * H], g[:, 2 * H:3 * H], g[:, 3 * H:]
        c = torch.sigmoid(f) * c + torch.sigmoid(i) * torch.tanh(gg)
        h = torch.sigmoid(o) * torch.tanh(c)
        out[:, t, :] = h
    return out, h


def lstm_stack(sd: Dict[str, Tensor], prefix: str, x: Tensor, layers: int, bidirectional: bool,
               quant: Quant = None, taps: Optional[dict] = None) -> Tuple[Tensor, Tensor]:
    """nn.LSTM(batch_first=True) restated; returns (outputs, last layer's final forward h)."""
    h_last = None
    for l in range(layers):
        outs = []
        for suffix, rev in (("", False),) + ((("_reverse", True),) if bidirectional else ()):
            o, h_last_dir = lstm_direction(
                x, sd[f"{prefix}.weight_ih_l{l}{suffix}"], sd[f"{prefix}.weight_hh_l{l}{suffix}"],
                sd[f"{prefix}.bias_ih_l{l}{suffix}"], sd[f"{prefix}.bias_hh_l{l}{suffix}"],
                reverse=rev, quant=quant, tag=f"{prefix}.l{l}{suffix}")
            outs.append(o)
            if not rev:
                h_last = h_last_dir
        x = torch.cat(outs, dim=-1) if bidirectional else outs[0]
        if taps is not None:
            taps[f"{prefix}.l{l}"] = x
    return x, h_last


def conv_bn_act(sd: Dict[str, Tensor], prefix: str, x: Tensor, act: str, quant: Quant = None,
                bn_stats: Optional[dict] = None) -> Tensor:
    """`bn_stats` (a dict) switches BatchNorm to training mode and receives `prefix -> (mean, biased var)`."""
    y = conv1d_k5(x, sd[f"{prefix}.0.conv.weight"], sd[f"{prefix}.0.conv.bias"], quant, prefix)
    if bn_stats is not None:
        y, m, v = batchnorm_train(y, sd[f"{prefix}.1.weight"], sd[f"{prefix}.1.bias"])
        bn_stats[prefix] = (m, v)
    else:
        y = batchnorm_eval(y, sd[f"{prefix}.1.weight"], sd[f"{prefix}.1.bias"],
                           sd[f"{prefix}.1.running_mean"], sd[f"{prefix}.1.running_var"])
    if act == "relu":
        return torch.clamp_min(y, 0.0)
    if act == "tanh":
        return torch.tanh(y)
    return y


# --------------------------------------------------------------------------------------------
# the path
# --------------------------------------------------------------------------------------------
def encoder_forward(sd: Dict[str, Tensor], x: Tensor, c_org: Tensor, freq: int = 32,
                    quant: Quant = None, taps: Optional[dict] = None, bn_stats: Optional[dict] = None) -> Tuple[List[Tensor], List[Tensor]]:
    """reference `encoder.py:53-97`: returns (codes_forward list, codes_backward list)."""
    B, _, T = x.shape
    h = torch.cat([x, c_org.unsqueeze(-1).expand(-1, -1, T)], dim=1)          # encoder.py:65-66
    for i in range(3):                                                          # encoder.py:69-70
        h = conv_bn_act(sd, f"encoder.convolutions.{i}", h, "relu", quant, bn_stats)
        if taps is not None:
            taps[f"encoder.conv{i}"] = h
    out, _ = lstm_stack(sd, "encoder.lstm", h.transpose(1, 2), 2, True, quant, taps)  # encoder.py:71-75
    dn = out.shape[-1] // 2
    out_f, out_b = out[:, :, :dn], out[:, :, dn:]                              # encoder.py:76-77
    codes_f = [out_f[:, i, :] for i in range(freq - 1, T, freq)]               # encoder.py:94
    codes_b = [out_b[:, i, :] for i in range(0, T, freq)]                      # encoder.py:95
    return codes_f, codes_b


def content_codes(codes_f: List[Tensor], codes_b: List[Tensor]) -> Tensor:
    """reference `model.py:99,143`: all forward codes, then all backward codes."""
    if len(codes_f) == 0:
        # the reference fails in torch.cat on the empty forward list when T < freq
        raise ValueError("T < freq: no forward content code exists (reference raises in torch.cat)")
    return torch.cat([torch.cat(codes_f, dim=-1), torch.cat(codes_b, dim=-1)], dim=-1)


def upsample_codes(codes_f: List[Tensor], codes_b: List[Tensor], T: int, freq: int) -> Tensor:
    """reference `model.py:115-123` -> (B, 2*dim_neck, T)."""
    B, dn = codes_b[0].shape
    up_f = torch.empty(B, dn, T, dtype=codes_b[0].dtype)
    up_b = torch.empty(B, dn, T, dtype=codes_b[0].dtype)
    nf = len(codes_f)
    for t in range(T):
        up_f[:, :, t] = codes_f[min(t // freq, nf - 1)]    # tail repeats the last forward code (:116-117)
        up_b[:, :, t] = codes_b[t // freq]                 # truncated to T (:119)
    return torch.cat([up_f, up_b], dim=1)


def decoder_forward(sd: Dict[str, Tensor], enc_out: Tensor, quant: Quant = None,
                    taps: Optional[dict] = None, bn_stats: Optional[dict] = None) -> Tensor:
    """reference `decoder.py:58-79`: (B,T,320) -> (B,T,80)."""
    h, _ = lstm_stack(sd, "decoder.lstm1", enc_out, 1, False, quant, taps)     # decoder.py:65
    h = h.transpose(1, 2)
    for i in range(3):                                                          # decoder.py:69-70
        h = conv_bn_act(sd, f"decoder.convolutions.{i}", h, "relu", quant, bn_stats)
        if taps is not None:
            taps[f"decoder.conv{i}"] = h
    h, _ = lstm_stack(sd, "decoder.lstm2", h.transpose(1, 2), 2, False, quant, taps)   # decoder.py:74
    w = sd["decoder.linear_projection.linear_layer.weight"]
    b = sd["decoder.linear_projection.linear_layer.bias"]
    return torch.matmul(_q(h, quant, "decoder.linear:x"), _q(w, quant, "decoder.linear:w").t()) + b   # decoder.py:77


def postnet_forward(sd: Dict[str, Tensor], mel: Tensor, quant: Quant = None, taps: Optional[dict] = None,
                    bn_stats: Optional[dict] = None) -> Tensor:
    """reference `postnet.py:50-63`."""
    h = mel
    for i in range(4):
        h = conv_bn_act(sd, f"postnet.convolutions.{i}", h, "tanh", quant, bn_stats)
        if taps is not None:
            taps[f"postnet.conv{i}"] = h
    return conv_bn_act(sd, "postnet.convolutions.4", h, "none", quant, bn_stats)


def generator_forward(sd: Dict[str, Tensor], x: Tensor, c_org: Tensor, c_trg: Optional[Tensor],
                      freq: int = 32, dtype: torch.dtype = torch.float32, quant: Quant = None,
                      taps: Optional[dict] = None, bn_stats: Optional[dict] = None):
    """reference `AutoEncoder.forward` (`model.py:68-146`) in eval mode; with `bn_stats` (a dict to fill) BatchNorm runs
    in training mode on batch statistics and the per-layer (mean, biased var) are recorded (see `batchnorm_train`).

    Returns (mel_outputs (B,80,T), mel_outputs_postnet (B,80,T), content_codes (B,dn*(nf+nb)))
    or only content_codes when `c_trg is None`."""
    sd = {k: (v.to(dtype) if v.is_floating_point() else v) for k, v in sd.items()}
    x, c_org = x.to(dtype), c_org.to(dtype)
    T = x.shape[-1]
    codes_f, codes_b = encoder_forward(sd, x, c_org, freq, quant, taps, bn_stats)
    codes = content_codes(codes_f, codes_b)
    if c_trg is None:
        return codes
    c_trg = c_trg.to(dtype)
    code_exp = upsample_codes(codes_f, codes_b, T, freq)
    enc_out = torch.cat([code_exp, c_trg.unsqueeze(-1).expand(-1, -1, T)], dim=1)   # model.py:124
    mel = decoder_forward(sd, enc_out.transpose(1, 2), quant, taps, bn_stats).transpose(2, 1)   # model.py:127
    mel_post = mel + postnet_forward(sd, mel, quant, taps, bn_stats)                            # model.py:131-132
    return mel, mel_post, codes


def batch_forward(sd: Dict[str, Tensor], batch: Tensor, c_org: Tensor, c_trg: Tensor,
                  overlap: float = 0.5, freq: int = 32, **kw) -> Tensor:
    """reference `AutoEncoder.batch_forward` (`model.py:395-437`): chunks of one utterance as a
    batch, then overlap-average with hop int(N*(1-overlap)); uncovered positions are ignored
    (the reference pads with NaN and takes nanmean)."""
    M, _, N = batch.shape
    _, post, _ = generator_forward(sd, batch, c_org.expand(M, -1), c_trg.expand(M, -1), freq, **kw)
    return overlap_average(post, overlap)


def overlap_average(post: Tensor, overlap: float = 0.5) -> Tensor:
    """(M,C,N) chunk outputs -> (C, N + (M-1)*hop), mean over the chunks covering each frame."""
    M, C, N = post.shape
    hop = int(N * (1 - overlap))
    total = N + (M - 1) * hop
    acc = torch.zeros(C, total, dtype=post.dtype)
    cnt = torch.zeros(total, dtype=post.dtype)
    for i in range(M):
        acc[:, i * hop:i * hop + N] += post[i]
        cnt[i * hop:i * hop + N] += 1
    return acc / cnt  # 0/0 -> nan where no chunk covers a frame, like nanmean of all-NaN


def speaker_encoder_forward(sd: Dict[str, Tensor], utterances: Tensor, quant: Quant = None) -> Tensor:
    """reference `SpeakerEncoder.forward` (`autovc/speaker_encoder/model.py:70-90`):
    LSTM(40,256,3) -> last layer's final h -> relu(Linear) -> L2 normalise.  (B,n_frames,40)->(B,256)."""
    layers = 1 + max(int(k.split("_l")[1][0]) for k in sd if k.startswith("lstm.weight_ih_l"))
    _, h_last = lstm_stack(sd, "lstm", utterances, layers, False, quant)
    e = torch.clamp_min(torch.matmul(_q(h_last, quant, "linear:x"), _q(sd["linear.weight"], quant, "linear:w").t()) + sd["linear.bias"], 0.0)
    return e / torch.linalg.norm(e, dim=1, keepdim=True)


def embed_utterance_from_partials(sd: Dict[str, Tensor], partials: Tensor) -> Tensor:
    """reference `embed_utterance` (`speaker_encoder/model.py:180-186`): mean of partial
    embeddings, re-normalised."""
    raw = speaker_encoder_forward(sd, partials).mean(dim=0)
    return raw / torch.linalg.norm(raw, 2)
