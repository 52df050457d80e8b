"""Deterministic synthetic weights and inputs for the AutoVC generator.

There are no pretrained checkpoints offline (reference `.gitignore:1-4` excludes `*.pt`), so
tests, `bench.py` and the golden-vector generator all build the *same* state dict from a seed
with numpy's PCG64 stream.  The names and shapes are the reference's state-dict contract
(`autovc/auto_encoder/net_layers.py:13,39` fix the `.conv.` / `.linear_layer.` infixes;
`encoder.py:39-51`, `decoder.py:37-56`, `postnet.py:19-48` fix the shapes).  Magnitudes follow
the reference initialisers (Xavier-uniform with relu/tanh/linear gain for conv/linear,
U(-1/sqrt(H), 1/sqrt(H)) for LSTM) and BatchNorm statistics are randomised so BN folding is
actually exercised (a fresh BatchNorm is the identity).
"""
from __future__ import annotations

import math
from typing import Dict, List, Tuple

import numpy as np
import torch

N_MELS = 80
LSTM2_HIDDEN = 1024
CONV_CH = 512
KSIZE = 5


def state_dict_spec(dim_neck: int = 32, dim_emb: int = 256, dim_pre: int = 512) -> List[Tuple[str, Tuple[int, ...], str]]:
    """Ordered (name, shape, kind) list of the generator's 100 state-dict entries."""
    spec: List[Tuple[str, Tuple[int, ...], str]] = []

    def conv_bn(prefix: str, cin: int, cout: int, gain: str):
        spec.append((f"{prefix}.0.conv.weight", (cout, cin, KSIZE), f"conv:{gain}"))
        spec.append((f"{prefix}.0.conv.bias", (cout,), f"bias:{cin * KSIZE}"))
        spec.append((f"{prefix}.1.weight", (cout,), "bn_gamma"))
        spec.append((f"{prefix}.1.bias", (cout,), "bn_beta"))
        spec.append((f"{prefix}.1.running_mean", (cout,), "bn_mean"))
        spec.append((f"{prefix}.1.running_var", (cout,), "bn_var"))
        spec.append((f"{prefix}.1.num_batches_tracked", (), "bn_count"))

    def lstm(prefix: str, cin: int, hidden: int, layers: int, bidir: bool):
        for l in range(layers):
            for suffix in ([""] + (["_reverse"] if bidir else [])):
                k = cin if l == 0 else hidden * (2 if bidir else 1)
                spec.append((f"{prefix}.weight_ih_l{l}{suffix}", (4 * hidden, k), f"lstm:{hidden}"))
                spec.append((f"{prefix}.weight_hh_l{l}{suffix}", (4 * hidden, hidden), f"lstm:{hidden}"))
                spec.append((f"{prefix}.bias_ih_l{l}{suffix}", (4 * hidden,), f"lstm:{hidden}"))
                spec.append((f"{prefix}.bias_hh_l{l}{suffix}", (4 * hidden,), f"lstm:{hidden}"))

    for i in range(3):
        conv_bn(f"encoder.convolutions.{i}", N_MELS + dim_emb if i == 0 else CONV_CH, CONV_CH, "relu")
    lstm("encoder.lstm", CONV_CH, dim_neck, 2, True)
    lstm("decoder.lstm1", 2 * dim_neck + dim_emb, dim_pre, 1, False)
    for i in range(3):
        conv_bn(f"decoder.convolutions.{i}", dim_pre, dim_pre, "relu")
    lstm("decoder.lstm2", dim_pre, LSTM2_HIDDEN, 2, False)
    spec.append(("decoder.linear_projection.linear_layer.weight", (N_MELS, LSTM2_HIDDEN), "linear:linear"))
    spec.append(("decoder.linear_projection.linear_layer.bias", (N_MELS,), f"bias:{LSTM2_HIDDEN}"))
    conv_bn("postnet.convolutions.0", N_MELS, CONV_CH, "tanh")
    for i in range(1, 4):
        conv_bn(f"postnet.convolutions.{i}", CONV_CH, CONV_CH, "tanh")
    conv_bn("postnet.convolutions.4", CONV_CH, N_MELS, "linear")
    return spec


_GAIN = {"relu": math.sqrt(2.0), "tanh": 5.0 / 3.0, "linear": 1.0}


def synth_state_dict(seed: int = 0, dim_neck: int = 32, dim_emb: int = 256, dim_pre: int = 512,
                     recurrent_gain: float = 1.0) -> Dict[str, torch.Tensor]:
    """Seeded state dict with the reference's names/shapes/dtypes.

    `recurrent_gain` scales the LSTM matrices; >1 drives the gates into their non-linear range
    (used by the stress fixtures)."""
    rng = np.random.default_rng(seed)
    sd: Dict[str, torch.Tensor] = {}
    for name, shape, kind in state_dict_spec(dim_neck, dim_emb, dim_pre):
        if kind.startswith("conv:") or kind.startswith("linear:"):
            gain = _GAIN[kind.split(":")[1]]
            rf = int(np.prod(shape[2:])) if len(shape) > 2 else 1
            fan_in, fan_out = shape[1] * rf, shape[0] * rf
            bound = gain * math.sqrt(6.0 / (fan_in + fan_out))
            arr = rng.uniform(-bound, bound, size=shape)
        elif kind.startswith("bias:"):
            bound = 1.0 / math.sqrt(int(kind.split(":")[1]))
            arr = rng.uniform(-bound, bound, size=shape)
        elif kind.startswith("lstm:"):
            bound = 1.0 / math.sqrt(int(kind.split(":")[1]))
            arr = rng.uniform(-bound, bound, size=shape)
            if len(shape) == 2:
                arr = arr * recurrent_gain
        elif kind == "bn_gamma":
            arr = rng.uniform(0.5, 1.5, size=shape)
        elif kind == "bn_beta":
            arr = rng.normal(0.0, 0.1, size=shape)
        elif kind == "bn_mean":
            arr = rng.normal(0.0, 0.1, size=shape)
        elif kind == "bn_var":
            arr = rng.uniform(0.5, 1.5, size=shape)
        elif kind == "bn_count":
            sd[name] = torch.tensor(1, dtype=torch.int64)
            continue
        else:  # pragma: no cover
            raise AssertionError(kind)
        sd[name] = torch.from_numpy(np.ascontiguousarray(arr.astype(np.float32)))
    return sd


def synth_inputs(batch: int, frames: int, seed: int = 1234, dim_emb: int = 256, same_speaker: bool = False):
    """x ~ U[0,1) (B,80,T) (the normalised-mel range, reference `audio/spectrogram.py:127`);
    embeddings = normalize(|N(0,1)|) (post-ReLU unit vectors, `speaker_encoder/model.py:85-88`)."""
    rng = np.random.default_rng(seed)
    x = rng.uniform(0.0, 1.0, size=(batch, N_MELS, frames)).astype(np.float32)

    def emb():
        e = np.abs(rng.normal(size=(batch, dim_emb))).astype(np.float32)
        return e / np.linalg.norm(e, axis=1, keepdims=True)

    c_org = emb()
    c_trg = c_org.copy() if same_speaker else emb()
    return torch.from_numpy(x), torch.from_numpy(c_org), torch.from_numpy(c_trg)


def synth_speaker_encoder(seed: int = 5, n_partials: int = 8, n_frames: int = 160):
    """Seeded d-vector speaker-encoder weights (reference names, `speaker_encoder/model.py:48-58`) and a batch of
    partial utterances (n_partials, n_frames, 40).  The draw order is the one `oracle/make_golden.py` used for
    `tests/golden/speaker_encoder_p8.npz` (seed 5, 8 partials of 160 frames)."""
    import numpy as np
    rng = np.random.default_rng(seed)
    names = []
    for l in range(3):
        names += [f"lstm.weight_ih_l{l}", f"lstm.weight_hh_l{l}", f"lstm.bias_ih_l{l}", f"lstm.bias_hh_l{l}"]
    names += ["linear.weight", "linear.bias"]
    shapes = {"lstm.weight_ih_l0": (1024, 40), "linear.weight": (256, 256), "linear.bias": (256,)}
    sd = {}
    for n in names:
        shp = shapes.get(n, (1024, 256) if "weight" in n else (1024,))
        sd[n] = torch.from_numpy(rng.uniform(-1 / 16, 1 / 16, size=shp).astype(np.float32))
    parts = torch.from_numpy(rng.normal(size=(n_partials, n_frames, 40)).astype(np.float32))
    return sd, parts


# ---- WaveRNN vocoder -------------------------------------------------------------------------------
WAVERNN = dict(hop_length=275, rnn_dims=512, res_out_dims=128, feat_dims=80, fc_dims=512, bits=9, upsample_factors=(5, 5, 11),
               compute_dims=128, pad=2, res_blocks=10, mode="MOL")                       # reference utils/hparams.py:92-107


def synth_wavernn_state_dict(seed: int = 11, logit_gain: float = 1.0) -> Dict[str, torch.Tensor]:
    """Seeded weights under the reference `WaveRNN` state-dict names (`autovc/wavernn/model.py:158-172`; the pretrained
    blob is absent, `.MISSING_LARGE_BLOBS:1`).  PyTorch default magnitudes (U(-1/sqrt(fan_in), ..) for conv / linear /
    GRU), BatchNorm statistics randomised so that folding is exercised, the box filters of the up-sampling convolutions
    (`model.py:81-82`: filled with 1/k) perturbed so that a wrong tap order shows."""
    rng = np.random.Generator(np.random.PCG64(seed))
    sd: Dict[str, torch.Tensor] = {}

    def uni(shape, fan_in, gain=1.0):
        b = gain / math.sqrt(fan_in)
        return torch.from_numpy(rng.uniform(-b, b, size=shape).astype(np.float32))

    def bn(prefix, c):
        sd[prefix + ".weight"] = torch.from_numpy(rng.uniform(0.5, 1.5, c).astype(np.float32))
        sd[prefix + ".bias"] = torch.from_numpy((0.1 * rng.standard_normal(c)).astype(np.float32))
        sd[prefix + ".running_mean"] = torch.from_numpy((0.1 * rng.standard_normal(c)).astype(np.float32))
        sd[prefix + ".running_var"] = torch.from_numpy(rng.uniform(0.5, 1.5, c).astype(np.float32))
        sd[prefix + ".num_batches_tracked"] = torch.tensor(1, dtype=torch.long)

    cd, feat, ro = WAVERNN["compute_dims"], WAVERNN["feat_dims"], WAVERNN["res_out_dims"]
    k = 2 * WAVERNN["pad"] + 1
    sd["upsample.resnet.conv_in.weight"] = uni((cd, feat, k), feat * k)
    bn("upsample.resnet.batch_norm", cd)
    for i in range(WAVERNN["res_blocks"]):
        p = f"upsample.resnet.layers.{i}"
        sd[p + ".conv1.weight"] = uni((cd, cd, 1), cd)
        sd[p + ".conv2.weight"] = uni((cd, cd, 1), cd)
        bn(p + ".batch_norm1", cd)
        bn(p + ".batch_norm2", cd)
    sd["upsample.resnet.conv_out.weight"] = uni((ro, cd, 1), cd)
    sd["upsample.resnet.conv_out.bias"] = uni((ro,), cd)
    for i, s in enumerate(WAVERNN["upsample_factors"]):
        ks = 2 * s + 1
        w = (1.0 / ks) * (1.0 + 0.2 * rng.uniform(-1, 1, size=(1, 1, 1, ks)))
        sd[f"upsample.up_layers.{2 * i + 1}.weight"] = torch.from_numpy(w.astype(np.float32))
    R, aux, fc = WAVERNN["rnn_dims"], ro // 4, WAVERNN["fc_dims"]
    sd["I.weight"] = uni((R, feat + aux + 1), feat + aux + 1)
    sd["I.bias"] = uni((R,), feat + aux + 1)
    for name, cin in (("rnn1", R), ("rnn2", R + aux)):
        sd[f"{name}.weight_ih_l0"] = uni((3 * R, cin), R)
        sd[f"{name}.weight_hh_l0"] = uni((3 * R, R), R)
        sd[f"{name}.bias_ih_l0"] = uni((3 * R,), R)
        sd[f"{name}.bias_hh_l0"] = uni((3 * R,), R)
    sd["fc1.weight"] = uni((fc, R + aux), R + aux)
    sd["fc1.bias"] = uni((fc,), R + aux)
    sd["fc2.weight"] = uni((fc, fc + aux), fc + aux)
    sd["fc2.bias"] = uni((fc,), fc + aux)
    # the 30 mixture logits: means in about +-0.5, log-scales around -3 (a trained vocoder's range), so that the
    # sample actually depends on the network instead of saturating the [-1, 1] clamp
    sd["fc3.weight"] = uni((30, fc), fc, gain=4.0 * logit_gain)
    b3 = uni((30,), fc)
    b3[20:] -= 3.0
    sd["fc3.bias"] = b3
    sd["step"] = torch.zeros(1, dtype=torch.long)
    return sd


def synth_uniforms(n_steps: int, batch: int, seed: int = 21) -> np.ndarray:
    """The random stream one `generate` consumes: per step and row, 10 uniforms for the mixture choice and one for the
    logistic (`autovc/wavernn/distribution.py:106,117`: both `uniform_(1e-5, 1 - 1e-5)`).  float32 (n_steps, batch, 11)."""
    rng = np.random.Generator(np.random.PCG64(seed))
    return rng.uniform(1e-5, 1.0 - 1e-5, size=(n_steps, batch, 11)).astype(np.float32)
