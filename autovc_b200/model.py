"""Drop-in for the reference generator `AutoEncoder` (reference `autovc/auto_encoder/model.py:27`, the class
upstream AutoVC calls `Generator`): same constructor defaults, same `forward(x, c_org, c_trg)` contract, same
`batch_forward`, same state-dict key names - but `forward` is ONE call into the hand-written sm_100a CUDA path
(`torch.ops.autovc_b200.generator_forward` -> C ABI `avc_forward`).  There is no PyTorch/CPU fallback.

Semantics pinned to the reference module in `.eval()` under `no_grad`: BatchNorm uses running
statistics and is folded into the convolution weights when the weights are handed to the C library.
With `batch_stat_bn=True` a module in `.train()` mode runs BatchNorm the way the reference's un-eval()'d module does
(batch statistics, running buffers updated with momentum 0.1): the forward pair of `AutoEncoder.learn`
(`model.py:306`, `:201`) and the literal behaviour of the reference's `convert`.  Forward only.

The sub-modules below exist only to own parameters under the reference's names
(`encoder.convolutions.0.0.conv.weight`, `decoder.linear_projection.linear_layer.weight`, ... - fixed by the
reference's `ConvNorm`/`LinearNorm` wrappers, `net_layers.py:13,39`) so that reference checkpoints load with
`load_state_dict`; they have no forward of their own.
"""
from __future__ import annotations

from typing import Dict, Optional, Tuple

import torch
import torch.nn as nn

from ._weights import PackedWeightsMixin
from .engine import Engine
from .synth import CONV_CH, KSIZE, LSTM2_HIDDEN, N_MELS


class _ParamOnly(nn.Module):
    def forward(self, *_a, **_k):  # pragma: no cover
        raise RuntimeError("parameter container: the generator runs as one fused CUDA op (AutoEncoder.forward); "
                           "autovc_b200 has no per-layer PyTorch path")


class ConvNorm(_ParamOnly):
    """Holds `conv.weight/bias` (reference `net_layers.py:23-49`; Xavier-uniform with gain)."""

    def __init__(self, cin: int, cout: int, w_init_gain: str = "linear"):
        super().__init__()
        self.conv = nn.Conv1d(cin, cout, kernel_size=KSIZE, stride=1, padding=2, dilation=1, bias=True)
        nn.init.xavier_uniform_(self.conv.weight, gain=nn.init.calculate_gain(w_init_gain))


class LinearNorm(_ParamOnly):
    """Holds `linear_layer.weight/bias` (reference `net_layers.py:3-20`)."""

    def __init__(self, cin: int, cout: int, w_init_gain: str = "linear"):
        super().__init__()
        self.linear_layer = nn.Linear(cin, cout, bias=True)
        nn.init.xavier_uniform_(self.linear_layer.weight, gain=nn.init.calculate_gain(w_init_gain))


def _conv_block(cin: int, cout: int, gain: str) -> nn.Sequential:
    return nn.Sequential(ConvNorm(cin, cout, gain), nn.BatchNorm1d(cout))


class Encoder(_ParamOnly):
    """Parameters of the content encoder (reference `encoder.py:22-51`)."""

    def __init__(self, dim_neck: int, dim_emb: int, freq: int):
        super().__init__()
        self.dim_neck, self.freq = dim_neck, freq
        self.convolutions = nn.ModuleList(
            [_conv_block(N_MELS + dim_emb if i == 0 else CONV_CH, CONV_CH, "relu") for i in range(3)])
        self.lstm = nn.LSTM(CONV_CH, dim_neck, 2, batch_first=True, bidirectional=True)


class Decoder(_ParamOnly):
    """Parameters of the decoder (reference `decoder.py:23-56`)."""

    def __init__(self, dim_neck: int, dim_emb: int, dim_pre: int):
        super().__init__()
        self.lstm1 = nn.LSTM(dim_neck * 2 + dim_emb, dim_pre, 1, batch_first=True)
        self.convolutions = nn.ModuleList([_conv_block(dim_pre, dim_pre, "relu") for _ in range(3)])
        self.lstm2 = nn.LSTM(dim_pre, LSTM2_HIDDEN, 2, batch_first=True)
        self.linear_projection = LinearNorm(LSTM2_HIDDEN, N_MELS)


class Postnet(_ParamOnly):
    """Parameters of the postnet (reference `postnet.py:16-48`)."""

    def __init__(self):
        super().__init__()
        blocks = [_conv_block(N_MELS, CONV_CH, "tanh")]
        blocks += [_conv_block(CONV_CH, CONV_CH, "tanh") for _ in range(3)]
        blocks += [_conv_block(CONV_CH, N_MELS, "linear")]
        self.convolutions = nn.ModuleList(blocks)


class AutoEncoder(PackedWeightsMixin, nn.Module):
    """B200-native AutoVC generator.

    Parameters mirror reference `AutoEncoder.__init__` (`model.py:31-38`; defaults from `utils/hparams.py:17-20`).
    `precision` is the only addition: "bf16" (tensor cores: bf16 operands, fp16 in the tanh-bounded postnet, fp32
    accumulate) or "fp32" (CUDA-core validation mode, <= 1e-4 of the reference).
    """

    def __init__(self, dim_neck: int = 32, dim_emb: int = 256, dim_pre: int = 512, freq: int = 32,
                 verbose: bool = True, device=None, precision: str = "bf16", batch_stat_bn: bool = False):
        super().__init__()
        if device is None:
            device = torch.device("cuda" if torch.cuda.is_available() else "cpu")
        self.verbose = verbose
        self.device = device if not isinstance(device, str) else torch.device(device)
        self.precision = precision
        self.batch_stat_bn = batch_stat_bn   # with .train(): BatchNorm uses batch statistics and updates its buffers
        self.logging = {}
        self.dims = (dim_neck, dim_emb, dim_pre, freq)
        self.encoder = Encoder(dim_neck, dim_emb, freq)
        self.decoder = Decoder(dim_neck, dim_emb, dim_pre)
        self.postnet = Postnet()
        self.train(False)  # parity contract: eval-mode BatchNorm
        self._init_tracking()
        self._engines: Dict[Tuple[int, str], Engine] = {}
        self._engine_keys: Dict[Tuple[int, str], Tuple[int, int]] = {}

    # -- weights (change tracking: _weights.PackedWeightsMixin) ----------------------------------------
    def flatten_params(self) -> torch.Tensor:
        """reference `AutoEncoder.flatten_params` (`model.py:379-383`): all parameters as one vector."""
        return torch.cat([p.data.view(-1) for p in self.parameters()], 0)

    def load_flattened_params(self, flattened_params: torch.Tensor) -> None:
        """reference `load_flattened_params` (`model.py:385-392`; the EMA weights of `learn`, `:275,321`).  Writes
        through `.data` like the reference and then invalidates the packed copy."""
        offset = 0
        for p in self.parameters():
            p.data.copy_(flattened_params[offset:offset + p.nelement()].view(p.size()))
            offset += p.nelement()
        self.invalidate()

    def _engine(self, device: torch.device, params_only: bool = False) -> Engine:
        idx = device.index if device.index is not None else torch.cuda.current_device()
        key = (idx, self.precision)
        eng = self._engines.get(key)
        if eng is None:
            eng = Engine(self.dims, idx, self.precision)
            self._engines[key] = eng
            self._engine_keys[key] = (None, None)
        wk = self._weights_key()
        if (self._engine_keys[key][1] != wk[1]) if params_only else (self._engine_keys[key] != wk):
            eng.load_state_dict(self.state_dict())
            self._engine_keys[key] = wk
        return eng

    # -- the path ----------------------------------------------------------------------------------
    def forward(self, x: torch.Tensor, c_org: torch.Tensor, c_trg: Optional[torch.Tensor], out=None):
        """reference `AutoEncoder.forward` (`model.py:68-146`).  `out` (extension): preallocated (mel, mel_post, codes) to fill.

        x (B,80,T) float, c_org/c_trg (B,dim_emb) -> (mel_outputs (B,80,T), mel_outputs_postnet (B,80,T),
        content_codes (B, dim_neck*(T//freq + ceil(T/freq)))); with `c_trg is None` only content_codes.
        Raises ValueError when T < freq, like the reference's torch.cat on an empty list."""
        if not x.is_cuda:
            raise RuntimeError("autovc_b200.AutoEncoder runs on CUDA (sm_100a) only; move the inputs to the GPU")
        if self.training and self.batch_stat_bn:
            return self._forward_bn_train(x, c_org, c_trg)
        eng = self._engine(x.device)
        c_org = c_org.to(x.device)
        if c_trg is None:
            return torch.ops.autovc_b200.generator_encode(x, c_org, eng.id)
        if out is not None:
            torch.ops.autovc_b200.generator_forward_out(x, c_org, c_trg.to(x.device), out[0], out[1], out[2], eng.id)
            return out
        mel, post, codes = torch.ops.autovc_b200.generator_forward(x, c_org, c_trg.to(x.device), eng.id)
        return mel, post, codes

    def _bn_modules(self):
        return [blk[1] for blk in list(self.encoder.convolutions) + list(self.decoder.convolutions) + list(self.postnet.convolutions)]

    def _forward_bn_train(self, x: torch.Tensor, c_org: torch.Tensor, c_trg: Optional[torch.Tensor]):
        """BatchNorm in training mode (reference: the module as constructed, `model.py:306,201`).  The kernels return
        each layer's batch mean / biased variance; the running buffers are updated here exactly like nn.BatchNorm1d
        (momentum 0.1, unbiased variance, num_batches_tracked += 1) - layers 0-2 only for the encode-only call."""
        eng = self._engine(x.device, params_only=True)
        c_org = c_org.to(x.device)
        if c_trg is None:
            codes, stats = torch.ops.autovc_b200.generator_encode_bn_train(x, c_org, eng.id)
            res, n_layers = codes, 3
        else:
            mel, post, codes, stats = torch.ops.autovc_b200.generator_forward_bn_train(x, c_org, c_trg.to(x.device), eng.id)
            res, n_layers = (mel, post, codes), 11
        n = x.shape[0] * x.shape[2]
        bns = self._bn_modules()[:n_layers]
        with torch.no_grad():
            st = stats.to(bns[0].running_mean.device)      # one transfer (the buffers live wherever the module lives)
            means = [st[i, 0, :bn.num_features] for i, bn in enumerate(bns)]
            uvars = [st[i, 1, :bn.num_features] for i, bn in enumerate(bns)]
            m = bns[0].momentum
            torch._foreach_mul_([bn.running_mean for bn in bns], 1 - m)
            torch._foreach_add_([bn.running_mean for bn in bns], means, alpha=m)
            torch._foreach_mul_([bn.running_var for bn in bns], 1 - m)
            torch._foreach_add_([bn.running_var for bn in bns], uvars, alpha=m * (n / max(n - 1, 1)))
            torch._foreach_add_([bn.num_batches_tracked for bn in bns], 1)
        return res

    def batch_forward(self, batch: torch.Tensor, c_org: torch.Tensor, c_trg: torch.Tensor, overlap: float = 0.5):
        """reference `AutoEncoder.batch_forward` (`model.py:395-437`): chunks of one utterance as a batch,
        overlap-averaged with hop int(N*(1-overlap))."""
        M = batch.size(0)
        _, output, _ = self(batch, c_org.expand(M, -1), c_trg.expand(M, -1))
        return overlap_average(output, overlap)

    def load(self, model_name: str, model_dir: str = "models/AutoVC", device=None):
        """reference `AutoEncoder.load` (`model.py:149-156`): checkpoint dict with key "model_state"."""
        self.device = device if device is not None else self.device
        path = model_dir.strip("/") + "/" + model_name
        checkpoint = torch.load(path, map_location="cpu")
        self.load_state_dict(checkpoint["model_state"])
        if self.verbose:
            print('Loaded auto encoder "%s" trained to step %s' % (path, checkpoint.get("step")))

    @property
    def last_launch_count(self) -> int:
        return max((e.last_launch_count for e in self._engines.values()), default=0)


def overlap_average(output: torch.Tensor, overlap: float = 0.5) -> torch.Tensor:
    """(M,C,N) -> (C, N+(M-1)*hop): mean over the chunks that cover each frame (reference pads with NaN and takes
    nanmean, `model.py:428-437`); frames no chunk covers come out NaN exactly like nanmean of all-NaN."""
    M, C, N = output.shape
    hop = int(N * (1 - overlap))
    total = N + (M - 1) * hop
    acc = torch.zeros(C, total, dtype=output.dtype, device=output.device)
    cnt = torch.zeros(total, dtype=output.dtype, device=output.device)
    for i in range(M):
        acc[:, i * hop:i * hop + N] += output[i]
        cnt[i * hop:i * hop + N] += 1
    return acc / cnt
