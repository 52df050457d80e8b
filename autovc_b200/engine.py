"""Host-side owner of one C-ABI context per (device, precision): weight hand-over, workspace cache and
the PyTorch custom op `autovc_b200::generator_forward` that calls avc_forward/avc_encode on the current
CUDA stream.  PyTorch is plumbing only here (device memory, streams)."""
from __future__ import annotations

import ctypes
import itertools
from typing import Dict, Optional, Tuple

import torch

from . import _lib as L

_ENGINES: Dict[int, "Engine"] = {}
_IDS = itertools.count(1)


class Engine:
    """One avc_ctx: a device, a precision and the packed weights of one state dict."""

    def __init__(self, dims: Tuple[int, int, int, int], device_index: int, precision: str = "bf16"):
        if precision not in ("bf16", "fp32"):
            raise ValueError("precision must be 'bf16' or 'fp32'")
        self.dims = tuple(int(d) for d in dims)
        self.device_index = int(device_index)
        self.precision = precision
        self._h = ctypes.c_void_p()
        d = L.AvcDims(*self.dims)
        rc = L.lib().avc_create(ctypes.byref(self._h), ctypes.byref(d), self.device_index,
                                L.AVC_PREC_BF16 if precision == "bf16" else L.AVC_PREC_FP32)
        L.raise_for(rc, None)
        self._ws: Optional[torch.Tensor] = None
        self.id = next(_IDS)
        _ENGINES[self.id] = self

    # -- weights -----------------------------------------------------------------------------
    def load_state_dict(self, sd: Dict[str, torch.Tensor], speaker_encoder: bool = False) -> None:
        """Hand a state dict to avc_load_weights (generator) or avc_se_load_weights (d-vector speaker encoder)."""
        keep, names = [], []
        for k, v in sd.items():
            if not v.is_floating_point():
                continue
            keep.append(v.detach().to(device="cpu", dtype=torch.float32).contiguous())
            names.append(k.encode())
        table = (L.AvcWeightDesc * len(keep))()
        for i, (n, t) in enumerate(zip(names, keep)):
            table[i].name = n
            table[i].data = ctypes.cast(t.data_ptr(), ctypes.POINTER(ctypes.c_float))
            table[i].numel = t.numel()
        load = L.lib().avc_se_load_weights if speaker_encoder else L.lib().avc_load_weights
        L.raise_for(load(self._h, table, len(keep)), self._h)

    def packed(self, which: str) -> torch.Tensor:
        n = L.lib().avc_debug_packed(self._h, which.encode(), None, 0)
        if n < 0:
            L.raise_for(int(n), self._h)
        out = torch.empty(int(n), dtype=torch.float32)
        L.lib().avc_debug_packed(self._h, which.encode(), ctypes.cast(out.data_ptr(), ctypes.POINTER(ctypes.c_float)), n)
        return out

    # -- run ---------------------------------------------------------------------------------
    def workspace_bytes(self, B: int, T: int) -> int:
        return int(L.lib().avc_workspace_bytes(self._h, B, T))

    def _workspace(self, B: int, T: int) -> torch.Tensor:
        need = self.workspace_bytes(B, T)
        if self._ws is None or self._ws.numel() < need:
            self._ws = None
            self._ws = torch.empty(need, dtype=torch.uint8, device=torch.device("cuda", self.device_index))
        return self._ws

    N_BN, BN_MAX_CH = 11, 512     # BatchNorm layers on the path (encoder 0-2, decoder 0-2, postnet 0-4), widest layer

    def forward(self, x: torch.Tensor, c_org: torch.Tensor, c_trg: Optional[torch.Tensor], out=None, bn_train: bool = False):
        """`out`: optional preallocated (mel, mel_post, codes) float32 CUDA tensors of the result shapes (steady-state
        pipelines that must not touch the allocator); default: fresh tensors, like the reference.
        `bn_train`: BatchNorm with batch statistics (avc_forward_bn_train); the result gains a trailing (11,2,512) tensor
        of per-layer batch mean / biased variance (rows of layers that did not run are NaN)."""
        B, n_mels, T = x.shape
        dn, E, _, freq = self.dims
        if n_mels != 80 or c_org.shape != (B, E) or (c_trg is not None and c_trg.shape != (B, E)):
            raise RuntimeError(f"shape mismatch: x {tuple(x.shape)}, c_org {tuple(c_org.shape)}, expected (B,80,T) and (B,{E})")
        if T < freq:
            # same condition avc_forward reports as AVC_ERR_T_TOO_SHORT; checked first so no allocation happens
            raise ValueError(f"T={T} < freq={freq}: no forward content code exists (the reference raises ValueError in torch.cat)")
        dev = x.device
        n_codes = dn * (T // freq + (T + freq - 1) // freq)
        codes = torch.empty(B, n_codes, dtype=torch.float32, device=dev)
        ws = self._workspace(B, T)
        stream = ctypes.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
        if bn_train:
            stats = torch.full((self.N_BN, 2, self.BN_MAX_CH), float("nan"), dtype=torch.float32, device=dev)
            if c_trg is None:
                mel = post = None
            elif out is not None:
                mel, post, codes = out
            else:
                mel = torch.empty(B, 80, T, dtype=torch.float32, device=dev)
                post = torch.empty(B, 80, T, dtype=torch.float32, device=dev)
            rc = L.lib().avc_forward_bn_train(self._h, x.data_ptr(), c_org.data_ptr(), c_trg.data_ptr() if c_trg is not None else None,
                                              B, T, mel.data_ptr() if mel is not None else None,
                                              post.data_ptr() if post is not None else None, codes.data_ptr(), stats.data_ptr(),
                                              ws.data_ptr(), ws.numel(), stream)
            L.raise_for(rc, self._h)
            return (codes, stats) if c_trg is None else (mel, post, codes, stats)
        if c_trg is None:
            rc = L.lib().avc_encode(self._h, x.data_ptr(), c_org.data_ptr(), B, T, codes.data_ptr(), ws.data_ptr(),
                                    ws.numel(), stream)
            L.raise_for(rc, self._h)
            return codes
        if out is not None:
            mel, post, codes = out
            for t, shp in ((mel, (B, 80, T)), (post, (B, 80, T)), (codes, (B, n_codes))):
                if tuple(t.shape) != shp or t.dtype != torch.float32 or not t.is_contiguous() or t.device != dev:
                    raise RuntimeError(f"out tensors must be contiguous float32 on {dev} with shapes (B,80,T), (B,80,T), (B,{n_codes})")
        else:
            mel = torch.empty(B, 80, T, dtype=torch.float32, device=dev)
            post = torch.empty(B, 80, T, dtype=torch.float32, device=dev)
        rc = L.lib().avc_forward(self._h, x.data_ptr(), c_org.data_ptr(), c_trg.data_ptr(), B, T, mel.data_ptr(),
                                 post.data_ptr(), codes.data_ptr(), ws.data_ptr(), ws.numel(), stream)
        L.raise_for(rc, self._h)
        return mel, post, codes

    # -- d-vector speaker encoder ----------------------------------------------------------------
    def se_forward(self, partials: torch.Tensor) -> torch.Tensor:
        """partials (N,T,40) float32 CUDA -> unit-norm embeddings (N,256) (avc_se_forward)."""
        N, T, C = partials.shape
        if C != 40:
            raise RuntimeError(f"speaker encoder expects (N, n_frames, 40) mel filterbanks, got {tuple(partials.shape)}")
        dev = partials.device
        out = torch.empty(N, 256, dtype=torch.float32, device=dev)
        if N == 0:
            return out
        if T == 0:
            raise ValueError("speaker encoder needs at least one frame")
        need = int(L.lib().avc_se_workspace_bytes(self._h, N, T))
        if self._ws is None or self._ws.numel() < need:
            self._ws = None
            self._ws = torch.empty(need, dtype=torch.uint8, device=torch.device("cuda", self.device_index))
        stream = ctypes.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
        rc = L.lib().avc_se_forward(self._h, partials.data_ptr(), N, T, out.data_ptr(), self._ws.data_ptr(),
                                    self._ws.numel(), stream)
        L.raise_for(rc, self._h)
        return out

    def se_reduce(self, embeds: torch.Tensor, offsets: torch.Tensor) -> torch.Tensor:
        """embeds (N,256) float32, offsets (S+1,) int32, both CUDA -> (S,256): mean of each row segment, renormalised."""
        S = offsets.numel() - 1
        out = torch.empty(S, 256, dtype=torch.float32, device=embeds.device)
        if S <= 0:
            return out
        stream = ctypes.c_void_p(torch.cuda.current_stream(embeds.device).cuda_stream)
        rc = L.lib().avc_se_reduce(self._h, embeds.data_ptr(), offsets.data_ptr(), S, out.data_ptr(), stream)
        L.raise_for(rc, self._h)
        return out

    _DT = {torch.float32: 0, torch.bfloat16: 1, torch.float16: 2}

    def debug_gemm(self, layer: str, A: torch.Tensor, T: int, act: int, out: Optional[torch.Tensor], use_tc: bool,
                   out_t: Optional[torch.Tensor] = None, res_t: Optional[torch.Tensor] = None) -> None:
        """Run one GEMM-shaped layer on caller buffers (tests): A (M, lda) channel-last rows."""
        M, lda = A.shape
        stream = ctypes.c_void_p(torch.cuda.current_stream(A.device).cuda_stream)
        rc = L.lib().avc_debug_gemm(self._h, layer.encode(), int(use_tc), A.data_ptr(), self._DT[A.dtype], lda, M, T, act,
                                    out.data_ptr() if out is not None else None,
                                    self._DT[out.dtype] if out is not None else 0,
                                    out.shape[1] if out is not None else 0,
                                    out_t.data_ptr() if out_t is not None else None,
                                    res_t.data_ptr() if res_t is not None else None, stream)
        L.raise_for(rc, self._h)

    PHASES = ("layout", "conv", "proj", "enc_lstm", "lstm1", "lstm2")

    def set_profiling(self, enable: bool) -> None:
        L.raise_for(L.lib().avc_set_profiling(self._h, int(enable)), self._h)

    def profile(self) -> Dict[str, float]:
        """Milliseconds per phase of the most recent profiled run (synchronises)."""
        buf = (ctypes.c_float * len(self.PHASES))()
        rc = L.lib().avc_get_profile(self._h, buf, len(self.PHASES))
        if rc < 0:
            L.raise_for(rc, self._h)
        return {k: float(buf[i]) for i, k in enumerate(self.PHASES)}

    @property
    def last_launch_count(self) -> int:
        return int(L.lib().avc_last_launch_count(self._h))

    @property
    def plan_builds(self) -> int:
        return int(L.lib().avc_plan_builds(self._h))

    @property
    def graph_replays(self) -> int:
        return int(L.lib().avc_graph_replays(self._h))

    def close(self) -> None:
        if self._h:
            L.lib().avc_destroy(self._h)
            self._h = ctypes.c_void_p()
        _ENGINES.pop(self.id, None)

    def __del__(self):  # pragma: no cover
        try:
            self.close()
        except Exception:
            pass


# ---------------------------------------------------------------------------------------------
# PyTorch custom op: the thin C-ABI exposed to torch programs
# ---------------------------------------------------------------------------------------------
_lib_def = torch.library.Library("autovc_b200", "DEF")
_lib_def.define("generator_forward(Tensor x, Tensor c_org, Tensor c_trg, int engine) -> (Tensor, Tensor, Tensor)")
_lib_def.define("generator_encode(Tensor x, Tensor c_org, int engine) -> Tensor")
_lib_def.define("speaker_embed(Tensor partials, int engine) -> Tensor")
_lib_def.define("generator_forward_bn_train(Tensor x, Tensor c_org, Tensor c_trg, int engine) -> (Tensor, Tensor, Tensor, Tensor)")
_lib_def.define("generator_encode_bn_train(Tensor x, Tensor c_org, int engine) -> (Tensor, Tensor)")
_lib_def.define("generator_forward_out(Tensor x, Tensor c_org, Tensor c_trg, Tensor(a!) mel, Tensor(b!) post, Tensor(c!) codes, int engine) -> ()")


def _prep(t: torch.Tensor) -> torch.Tensor:
    return t.detach().to(dtype=torch.float32).contiguous()


def _forward_cuda(x, c_org, c_trg, engine: int):
    return _ENGINES[engine].forward(_prep(x), _prep(c_org), _prep(c_trg))


def _encode_cuda(x, c_org, engine: int):
    return _ENGINES[engine].forward(_prep(x), _prep(c_org), None)


def _forward_bn_train_cuda(x, c_org, c_trg, engine: int):
    return _ENGINES[engine].forward(_prep(x), _prep(c_org), _prep(c_trg), bn_train=True)


def _encode_bn_train_cuda(x, c_org, engine: int):
    return _ENGINES[engine].forward(_prep(x), _prep(c_org), None, bn_train=True)


def _forward_out_cuda(x, c_org, c_trg, mel, post, codes, engine: int):
    _ENGINES[engine].forward(_prep(x), _prep(c_org), _prep(c_trg), out=(mel, post, codes))


def _speaker_embed_cuda(partials, engine: int):
    return _ENGINES[engine].se_forward(_prep(partials))


def _no_cpu(*_a, **_k):
    raise RuntimeError("autovc_b200 ops have only a CUDA (sm_100a) implementation; there is no CPU fallback")


_lib_impl_cuda = torch.library.Library("autovc_b200", "IMPL", "CUDA")
_lib_impl_cuda.impl("generator_forward", _forward_cuda)
_lib_impl_cuda.impl("generator_encode", _encode_cuda)
_lib_impl_cuda.impl("speaker_embed", _speaker_embed_cuda)
_lib_impl_cuda.impl("generator_forward_out", _forward_out_cuda)
_lib_impl_cuda.impl("generator_forward_bn_train", _forward_bn_train_cuda)
_lib_impl_cuda.impl("generator_encode_bn_train", _encode_bn_train_cuda)
_lib_impl_cpu = torch.library.Library("autovc_b200", "IMPL", "CPU")
_lib_impl_cpu.impl("generator_forward", _no_cpu)
_lib_impl_cpu.impl("generator_encode", _no_cpu)
_lib_impl_cpu.impl("speaker_embed", _no_cpu)
_lib_impl_cpu.impl("generator_forward_out", _no_cpu)
_lib_impl_cpu.impl("generator_forward_bn_train", _no_cpu)
_lib_impl_cpu.impl("generator_encode_bn_train", _no_cpu)
