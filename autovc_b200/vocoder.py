"""Drop-in for the inference side of the reference `WaveRNN` vocoder (`autovc/wavernn/model.py:95-460`): mel -> waveform.

Same constructor hyper-parameters (`utils/hparams.py:92-107`), same state-dict key names (`I.weight`, `rnn1.weight_ih_l0`,
`upsample.resnet.layers.0.conv1.weight`, `upsample.up_layers.1.weight`, ...), same `generate(melspectrogram, batched,
target, overlap, mu_law)` contract and return value (float32 numpy waveform of `(T - 1) * 275` samples) - but the
conditioning network and the whole per-sample loop are hand-written sm_100a CUDA (`csrc/wavernn.cu`, C ABI `avc_voc_*`):
one persistent weight-stationary kernel runs all `target + 2 * overlap` steps of all folds.  There is no PyTorch / CPU
fallback.

What stays on the host, as in the reference: the fold arithmetic (`fold_with_overlap`, :358-401, here a table of row
windows handed to the kernel) and the float64 cross-fade / fade-out on the returned numpy array (:403-456, :318-327).

Beyond the reference: `generate_batch` runs the folds of MANY utterances as one batch (the reference folds one utterance
per call), and the random stream is explicit - `uniforms=` for reproducible / parity runs, otherwise a Philox stream
keyed by `seed` (the reference draws from torch's global generator).

Training (`forward`, :175-211) and the 'RAW' softmax mode are out of scope.
"""
from __future__ import annotations

import ctypes
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np
import torch
import torch.nn as nn

from . import _lib as L
from ._weights import PackedWeightsMixin
from .synth import WAVERNN

HOP = WAVERNN["hop_length"]


def fold_params(total_len: int, target: int, overlap: int) -> Tuple[int, int]:
    """`fold_with_overlap` (model.py:381-391): (number of folds, zero padding appended to the conditioning)."""
    num_folds = (total_len - overlap) // (target + overlap)
    extended_len = num_folds * (overlap + target) + overlap
    remaining = total_len - extended_len
    padding = 0
    if remaining != 0:
        num_folds += 1
        padding = target + 2 * overlap - remaining
    return num_folds, padding


def xfade_and_unfold(y: np.ndarray, target: int, overlap: int) -> np.ndarray:
    """`xfade_and_unfold` (model.py:403-456): float64 (num_folds, target + 2*overlap), modified in place -> 1-D float64."""
    num_folds, length = y.shape
    target = length - 2 * overlap
    total_len = num_folds * (target + overlap) + overlap
    silence_len = overlap // 2
    fade_len = overlap - silence_len
    t = np.linspace(-1, 1, fade_len, dtype=np.float64)
    fade_in = np.concatenate([np.zeros(silence_len, dtype=np.float64), np.sqrt(0.5 * (1 + t))])
    fade_out = np.concatenate([np.ones(silence_len, dtype=np.float64), np.sqrt(0.5 * (1 - t))])
    y[:, :overlap] *= fade_in
    y[:, -overlap:] *= fade_out
    unfolded = np.zeros(total_len, dtype=np.float64)
    for i in range(num_folds):
        start = i * (target + overlap)
        unfolded[start:start + target + 2 * overlap] += y[i]
    return unfolded


class _ParamsOnly(nn.Module):
    def __init__(self, **params):
        super().__init__()
        for k, v in params.items():
            self.register_parameter(k, nn.Parameter(v, requires_grad=False))


class WaveRNN(PackedWeightsMixin, nn.Module):
    _skip_in_key = ("num_batches_tracked", "step")

    def __init__(self, hop_length=WAVERNN["hop_length"], rnn_dims=WAVERNN["rnn_dims"], res_out_dims=WAVERNN["res_out_dims"],
                 feat_dims=WAVERNN["feat_dims"], fc_dims=WAVERNN["fc_dims"], bits=WAVERNN["bits"],
                 upsample_factors=WAVERNN["upsample_factors"], compute_dims=WAVERNN["compute_dims"], pad=WAVERNN["pad"],
                 res_blocks=WAVERNN["res_blocks"], mode=WAVERNN["mode"], device=None):
        super().__init__()
        given = dict(hop_length=hop_length, rnn_dims=rnn_dims, res_out_dims=res_out_dims, feat_dims=feat_dims, fc_dims=fc_dims,
                     upsample_factors=tuple(upsample_factors), compute_dims=compute_dims, pad=pad, res_blocks=res_blocks, mode=mode)
        fixed = {k: WAVERNN[k] for k in given}
        if given != fixed:
            raise ValueError(f"the CUDA path is built for the reference configuration {fixed}")
        if device is None:
            device = torch.device("cuda" if torch.cuda.is_available() else "cpu")
        self.device = torch.device(device) if isinstance(device, str) else device
        self.hop_length, self.rnn_dims, self.fc_dims, self.pad, self.mode, self.bits = hop_length, rnn_dims, fc_dims, pad, mode, bits
        self.aux_dims = res_out_dims // 4
        self.n_classes = 30
        # parameter containers under the reference's names (model.py:158-172); generate() is the fused CUDA path
        aux = self.aux_dims
        self.upsample = nn.Module()
        self.upsample.resnet = nn.Module()
        k = 2 * pad + 1
        self.upsample.resnet.conv_in = nn.Conv1d(feat_dims, compute_dims, kernel_size=k, bias=False)
        self.upsample.resnet.batch_norm = nn.BatchNorm1d(compute_dims)
        self.upsample.resnet.layers = nn.ModuleList()
        for _ in range(res_blocks):
            blk = nn.Module()
            blk.conv1 = nn.Conv1d(compute_dims, compute_dims, kernel_size=1, bias=False)
            blk.conv2 = nn.Conv1d(compute_dims, compute_dims, kernel_size=1, bias=False)
            blk.batch_norm1 = nn.BatchNorm1d(compute_dims)
            blk.batch_norm2 = nn.BatchNorm1d(compute_dims)
            self.upsample.resnet.layers.append(blk)
        self.upsample.resnet.conv_out = nn.Conv1d(compute_dims, res_out_dims, kernel_size=1)
        self.upsample.up_layers = nn.ModuleList()
        for s in upsample_factors:
            conv = nn.Conv2d(1, 1, kernel_size=(1, 2 * s + 1), padding=(0, s), bias=False)
            conv.weight.data.fill_(1.0 / (2 * s + 1))
            self.upsample.up_layers.append(nn.Identity())          # the parameter-free Stretch2d keeps the indices 1, 3, 5
            self.upsample.up_layers.append(conv)
        self.I = nn.Linear(feat_dims + aux + 1, rnn_dims)
        self.rnn1 = nn.GRU(rnn_dims, rnn_dims, batch_first=True)
        self.rnn2 = nn.GRU(rnn_dims + aux, rnn_dims, batch_first=True)
        self.fc1 = nn.Linear(rnn_dims + aux, fc_dims)
        self.fc2 = nn.Linear(fc_dims + aux, fc_dims)
        self.fc3 = nn.Linear(fc_dims, self.n_classes)
        self.register_buffer("step", torch.zeros(1, dtype=torch.long))
        self.train(False)
        for p in self.parameters():
            p.requires_grad_(False)
        self._init_tracking()
        self._handles: Dict[int, ctypes.c_void_p] = {}
        self._keys: Dict[int, object] = {}
        self.last_launch_count = 0
        self.last_rows = 0
        self.last_steps = 0

    # -- weights -----------------------------------------------------------------------------------
    def _handle(self, device: torch.device) -> ctypes.c_void_p:
        idx = device.index if device.index is not None else torch.cuda.current_device()
        h = self._handles.get(idx)
        if h is None:
            h = ctypes.c_void_p()
            rc = L.lib().avc_voc_create(ctypes.byref(h), idx)
            if rc != L.AVC_OK:
                raise RuntimeError(f"avc_voc_create failed ({rc}): {L.lib().avc_voc_last_error(None).decode()}")
            self._handles[idx] = h
            self._keys[idx] = None
        wk = self._weights_key()[0]
        if self._keys[idx] != wk:
            sd = {k: v.detach().to("cpu", torch.float32).contiguous() for k, v in self.state_dict().items()
                  if not k.endswith("num_batches_tracked") and k != "step"}
            table = (L.AvcWeightDesc * len(sd))()
            for i, (k, v) in enumerate(sd.items()):
                table[i] = L.AvcWeightDesc(k.encode(), ctypes.cast(v.data_ptr(), ctypes.POINTER(ctypes.c_float)), v.numel())
            rc = L.lib().avc_voc_load_weights(h, table, len(sd))
            if rc != L.AVC_OK:
                msg = L.lib().avc_voc_last_error(h).decode()
                raise (KeyError if rc == L.AVC_ERR_MISSING_WEIGHT else RuntimeError)(f"avc_voc_load_weights failed ({rc}): {msg}")
            self._keys[idx] = wk
        return h

    def load(self, model_name: str, model_dir: str = "models/WaveRNN", device=None):
        """reference `WaveRNN.load` shape of use: a checkpoint that is either the state dict or has it under "model_state"."""
        self.device = device if device is not None else self.device
        ckpt = torch.load(model_dir.strip("/") + "/" + model_name, map_location="cpu")
        self.load_state_dict(ckpt.get("model_state", ckpt) if isinstance(ckpt, dict) else ckpt, strict=False)

    def _check(self, rc: int, h, what: str):
        if rc != L.AVC_OK:
            msg = L.lib().avc_voc_last_error(h).decode()
            raise (ValueError if rc in (L.AVC_ERR_INVALID_ARG, L.AVC_ERR_UNSUPPORTED) else RuntimeError)(f"{what} failed ({rc}): {msg}")

    def _dev(self) -> torch.device:
        if self.device.type != "cuda":
            if not torch.cuda.is_available():
                raise RuntimeError("autovc_b200.WaveRNN runs on CUDA (sm_100a) only; there is no CPU fallback")
            return torch.device("cuda")
        return self.device

    # -- the path ----------------------------------------------------------------------------------
    def condition(self, mels: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
        """`UpsampleNetwork.forward` on the zero-padded mel (model.py:252-253): mels (U, 80, T) CUDA float32 ->
        (m_up (U, T*275, 80), aux (U, T, 128) at frame rate; the reference's aux is this repeated 275x along time)."""
        if not mels.is_cuda:
            raise RuntimeError("autovc_b200.WaveRNN runs on CUDA (sm_100a) only; move the mel to the GPU")
        mels = mels.detach().to(torch.float32).contiguous()
        U, F, T = mels.shape
        if F != 80:
            raise ValueError("mel must be (U, 80, T)")
        h = self._handle(mels.device)
        m_up = torch.empty((U, T * HOP, 80), dtype=torch.float32, device=mels.device)
        aux = torch.empty((U, T, 128), dtype=torch.float32, device=mels.device)
        ws = torch.empty(int(L.lib().avc_voc_condition_workspace_bytes(h, U, T)), dtype=torch.uint8, device=mels.device)
        stream = ctypes.c_void_p(torch.cuda.current_stream(mels.device).cuda_stream)
        self._check(L.lib().avc_voc_condition(h, mels.data_ptr(), U, T, m_up.data_ptr(), aux.data_ptr(), ws.data_ptr(), stream), h, "avc_voc_condition")
        self.last_launch_count = int(L.lib().avc_voc_last_launch_count(h))
        return m_up, aux

    def sample_rows(self, m_up: torch.Tensor, aux: torch.Tensor, row_utt, row_start, steps: int, uniforms=None, seed: int = 0,
                    forced_x=None, return_logits: bool = False):
        """The per-sample loop (model.py:265-310) on rows = windows (utterance, start) of `steps` samples.
        -> samples (rows, steps) CUDA float32 [, logits (steps, rows, 30)].  uniforms: (steps, rows, 11) or None (Philox, seed)."""
        dev = m_up.device
        h = self._handle(dev)
        U, Ltot, _ = m_up.shape
        T = aux.shape[1]
        row_utt = torch.as_tensor(np.asarray(row_utt, dtype=np.int32)).to(dev)
        row_start = torch.as_tensor(np.asarray(row_start, dtype=np.int32)).to(dev)
        R = int(row_utt.numel())
        maxr = int(L.lib().avc_voc_max_rows())
        samples = torch.empty((R, steps), dtype=torch.float32, device=dev)
        logits = torch.empty((steps, R, 30), dtype=torch.float32, device=dev) if return_logits else None
        if uniforms is not None:
            uniforms = torch.as_tensor(uniforms, dtype=torch.float32).to(dev).contiguous()
            if tuple(uniforms.shape) != (steps, R, 11):
                raise ValueError(f"uniforms must be (steps, rows, 11) = {(steps, R, 11)}, got {tuple(uniforms.shape)}")
        if forced_x is not None:
            forced_x = torch.as_tensor(forced_x, dtype=torch.float32).to(dev).contiguous()
            if tuple(forced_x.shape) != (steps, R):
                raise ValueError(f"forced_x must be (steps, rows) = {(steps, R)}")
        stream = ctypes.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
        launches = 0
        for r0 in range(0, R, maxr):                                   # at most avc_voc_max_rows() rows per launch
            n = min(maxr, R - r0)
            ws = torch.empty(int(L.lib().avc_voc_generate_workspace_bytes(h, n, steps)), dtype=torch.uint8, device=dev)
            u = uniforms[:, r0:r0 + n].contiguous() if uniforms is not None else None
            f = forced_x[:, r0:r0 + n].contiguous() if forced_x is not None else None
            lg = torch.empty((steps, n, 30), dtype=torch.float32, device=dev) if return_logits else None
            out = samples[r0:r0 + n]
            self._check(L.lib().avc_voc_generate(
                h, m_up.data_ptr(), aux.data_ptr(), U, T, row_utt[r0:].data_ptr(), row_start[r0:].data_ptr(), n, steps,
                u.data_ptr() if u is not None else None, ctypes.c_uint64((int(seed) + r0) & 0xFFFFFFFFFFFFFFFF),
                f.data_ptr() if f is not None else None, out.data_ptr(), lg.data_ptr() if lg is not None else None, ws.data_ptr(), stream),
                h, "avc_voc_generate")
            launches += int(L.lib().avc_voc_last_launch_count(h))
            if return_logits:
                logits[:, r0:r0 + n] = lg
        self.last_launch_count = launches
        self.last_rows, self.last_steps = R, steps
        return (samples, logits) if return_logits else samples

    def generate_batch(self, mels: Sequence, batched: bool = True, target: int = 11_000, overlap: int = 550, uniforms=None,
                       seed: int = 0) -> List[np.ndarray]:
        """`generate` for several utterances of the SAME number of frames at once: their folds form one batch of rows."""
        dev = self._dev()
        ms = [torch.as_tensor(m, dtype=torch.float32) for m in mels]
        ms = [m[0] if m.dim() == 3 else m for m in ms]
        if len({tuple(m.shape) for m in ms}) != 1:
            raise ValueError("generate_batch needs mels of one shape (80, T); group the work list by length")
        T = ms[0].shape[-1]
        wave_len = (T - 1) * HOP
        if wave_len < 20 * HOP:
            raise ValueError("mel too short: the reference's 20-frame fade-out needs at least 21 frames (model.py:327)")
        m_up, aux = self.condition(torch.stack(ms).to(dev))
        total = T * HOP
        if batched:
            nf, _ = fold_params(total, target, overlap)
            steps = target + 2 * overlap
            starts = [i * (target + overlap) for i in range(nf)]
        else:
            nf, steps, starts = 1, total, [0]
        row_utt = np.repeat(np.arange(len(ms), dtype=np.int32), nf)
        row_start = np.tile(np.asarray(starts, dtype=np.int32), len(ms))
        samples = self.sample_rows(m_up, aux, row_utt, row_start, steps, uniforms=uniforms, seed=seed)
        out = samples.cpu().numpy().astype(np.float64).reshape(len(ms), nf, steps)
        waves = []
        for i in range(len(ms)):
            y = xfade_and_unfold(out[i], target, overlap) if batched else out[i, 0]
            y = y[:wave_len]
            y[-20 * HOP:] *= np.linspace(1, 0, 20 * HOP)
            waves.append(y.astype(np.float32))
        return waves

    def generate(self, melspectrogram, batched: bool = True, target: int = 11_000, overlap: int = 550, mu_law: bool = False,
                 uniforms=None, seed: int = 0) -> np.ndarray:
        """reference `WaveRNN.generate` (model.py:214-329): melspectrogram (1, 80, T) -> float32 waveform ((T-1)*275,)."""
        return self.generate_batch([melspectrogram], batched, target, overlap, uniforms, seed)[0]

    def forward(self, x, mels):
        raise NotImplementedError("training forward of the vocoder is out of scope (SURVEY.md 8f); use generate()")

    def __del__(self):  # pragma: no cover
        try:
            for h in self._handles.values():
                L.lib().avc_voc_destroy(h)
        except Exception:
            pass
