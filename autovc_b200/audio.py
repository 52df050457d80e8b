"""Mel front-end on the device: drop-in for the spectrogram functions of the reference
(`autovc/audio/spectrogram.py`) that produce the generator's `x` and the speaker encoder's partial utterances.

    mel_spec_auto_encoder(wav, cut=False, ...)      reference :62-142   -> (80, T) tensor, or a list of (80, 400) chunks
    mel_spec_speaker_encoder(wav, cut=False, ...)   reference :144-217  -> (T, 40), or a list of (160, 40) partials
    compute_partial_slices(...)                     reference :248-311  (host integer arithmetic, unchanged contract)
    amp_to_db / db_to_amp / normalize_spec / denormalize_spec           (:14-60, torch)

The STFT + mel projection + dB/normalise run as ONE CUDA kernel per call (`csrc/mel_frontend.cu`, C ABI `avc_mel_*`),
with librosa's semantics (centre padding, periodic Hann zero-padded to n_fft, float64 transform, Slaney filterbank).
Outputs stay on the device: they feed `AutoEncoder.forward` / `SpeakerEncoder.forward` without a host round trip (the
reference returns CPU tensors / numpy arrays; call `.cpu()` where that is needed).  There is no CPU fallback.
Wav loading, resampling, VAD and noise reduction (`autovc/audio/tools.py`) stay on the caller's side.
"""
from __future__ import annotations

import ctypes
from typing import Dict, List, Optional, Tuple

import numpy as np
import torch

from . import _lib as L

AE_SPEC = dict(sr=22050, n_mels=80, n_fft=2048, hop_length=275, window_length=1100, fmin=40,
               mel_window_step=12.5, partial_utterance_n_frames=400)          # reference utils/hparams.py:6-15
SE_SPEC = dict(sr=16000, n_mels=40, mel_window_step=10, mel_window_length=25, partial_utterance_n_frames=160)   # :52-58


class MelFrontend:
    """One `avc_mel` handle: a fixed (sr, n_fft, hop, window, filterbank, output transform) on one device."""

    def __init__(self, sr: int, n_fft: int, hop_length: int, win_length: Optional[int], n_mels: int, fmin: float = 0.0,
                 fmax: Optional[float] = None, power: int = 1, log_normalize: bool = False, min_level_db: float = -100.0,
                 pad_mode: str = "reflect", frames_major: bool = False, device_index: int = 0):
        if pad_mode not in ("reflect", "constant"):
            raise ValueError("pad_mode must be 'reflect' (librosa <= 0.9, the reference's vintage) or 'constant'")
        self.cfg = L.AvcMelConfig(int(sr), int(n_fft), int(hop_length), int(win_length or 0), int(n_mels), float(fmin),
                                  float(fmax or 0.0), int(power), int(bool(log_normalize)), float(min_level_db),
                                  int(pad_mode == "reflect"), int(bool(frames_major)))
        self.device_index = int(device_index)
        self._h = ctypes.c_void_p()
        rc = L.lib().avc_mel_create(ctypes.byref(self._h), ctypes.byref(self.cfg), self.device_index)
        if rc != L.AVC_OK:
            raise RuntimeError(f"avc_mel_create failed ({rc}): {L.lib().avc_mel_last_error(None).decode()}")

    def num_frames(self, n_samples: int) -> int:
        return int(L.lib().avc_mel_num_frames(self._h, int(n_samples)))

    def filterbank(self) -> torch.Tensor:
        n = int(L.lib().avc_mel_filterbank(self._h, None, 0))
        out = torch.empty(n, dtype=torch.float32)
        L.lib().avc_mel_filterbank(self._h, ctypes.cast(out.data_ptr(), ctypes.POINTER(ctypes.c_float)), n)
        return out.view(self.cfg.n_mels, 1 + self.cfg.n_fft // 2)

    def __call__(self, wav: torch.Tensor) -> torch.Tensor:
        """wav (n,) or (B, n) float32 CUDA -> (n_mels, T) / (B, n_mels, T), or (T, n_mels) / (B, T, n_mels) when frames_major."""
        if not wav.is_cuda:
            raise RuntimeError("autovc_b200 mel front-end runs on CUDA (sm_100a) only; move the waveform to the GPU")
        single = wav.dim() == 1
        w = wav.detach().to(torch.float32)
        w = (w.unsqueeze(0) if single else w).contiguous()
        B, n = w.shape
        T = self.num_frames(n)
        shape = (B, T, self.cfg.n_mels) if self.cfg.frames_major else (B, self.cfg.n_mels, T)
        out = torch.empty(shape, dtype=torch.float32, device=w.device)
        stream = ctypes.c_void_p(torch.cuda.current_stream(w.device).cuda_stream)
        rc = L.lib().avc_mel_forward(self._h, w.data_ptr(), B, n, w.stride(0), out.data_ptr(), stream)
        if rc != L.AVC_OK:
            msg = L.lib().avc_mel_last_error(self._h).decode()
            raise (ValueError if rc == L.AVC_ERR_INVALID_ARG else RuntimeError)(f"avc_mel_forward failed ({rc}): {msg}")
        return out[0] if single else out

    def close(self) -> None:
        if self._h:
            L.lib().avc_mel_destroy(self._h)
            self._h = ctypes.c_void_p()

    def __del__(self):  # pragma: no cover
        try:
            self.close()
        except Exception:
            pass


_FRONTENDS: Dict[Tuple, MelFrontend] = {}


def _frontend(device: torch.device, **kw) -> MelFrontend:
    idx = device.index if device.index is not None else torch.cuda.current_device()
    key = (idx,) + tuple(sorted(kw.items()))
    fe = _FRONTENDS.get(key)
    if fe is None:
        fe = _FRONTENDS[key] = MelFrontend(device_index=idx, **kw)
    return fe


def _to_device(wav, device) -> torch.Tensor:
    if isinstance(wav, torch.Tensor):
        t = wav
    else:
        t = torch.from_numpy(np.ascontiguousarray(wav, dtype=np.float32))
    if not t.is_cuda:
        t = t.to(device if device is not None else "cuda")
    return t.to(torch.float32)


# ---- reference autovc/audio/spectrogram.py:14-60 ----------------------------------------------------------------------
def normalize_spec(spectrogram: torch.Tensor, min_level_db: float = -100) -> torch.Tensor:
    return torch.clamp((spectrogram - min_level_db) / -min_level_db, 0, 1)


def denormalize_spec(spectrogram: torch.Tensor, min_level_db: float = -100) -> torch.Tensor:
    return torch.clamp(spectrogram, 0, 1) * -min_level_db + min_level_db


def amp_to_db(amplitude: torch.Tensor) -> torch.Tensor:
    return 20 * torch.log10(torch.clamp(amplitude, min=1e-5))


def db_to_amp(power: torch.Tensor) -> torch.Tensor:
    return torch.pow(10.0, power * 0.05)


def compute_partial_slices(n_samples: int, sr: int, partial_utterance_n_frames: int = 160, min_pad_coverage: float = 0.75,
                           overlap: float = 0.5, mel_window_step: float = 10):
    """reference `compute_partial_slices` (:248-311): where to cut a waveform / its mel spectrogram into partial
    utterances.  Same integer arithmetic, same return value ([slice] of samples, [slice] of frames)."""
    assert 0 <= overlap < 1
    assert 0 < min_pad_coverage <= 1
    samples_per_frame = int((sr * mel_window_step / 1000))
    n_frames = int(np.ceil((n_samples + 1) / samples_per_frame))
    frame_step = max(int(np.round(partial_utterance_n_frames * (1 - overlap))), 1)
    wav_slices, mel_slices = [], []
    steps = max(1, n_frames - partial_utterance_n_frames + frame_step + 1)
    for i in range(0, steps, frame_step):
        mel_slices.append(slice(i, i + partial_utterance_n_frames))
        wav_slices.append(slice(i * samples_per_frame, (i + partial_utterance_n_frames) * samples_per_frame))
    last = wav_slices[-1]
    coverage = (n_samples - last.start) / (last.stop - last.start)
    if coverage < min_pad_coverage and len(mel_slices) > 1:
        mel_slices, wav_slices = mel_slices[:-1], wav_slices[:-1]
    return wav_slices, mel_slices


def _pad_to(wav: torch.Tensor, length: int) -> torch.Tensor:
    return torch.nn.functional.pad(wav, (0, length - wav.shape[-1])) if length > wav.shape[-1] else wav


def mel_spec_auto_encoder(wav, sr: int = AE_SPEC["sr"], cut: bool = False, n_mels: int = AE_SPEC["n_mels"],
                          n_fft: int = AE_SPEC["n_fft"], hop_length: int = AE_SPEC["hop_length"],
                          window_length: int = AE_SPEC["window_length"], fmin: float = AE_SPEC["fmin"], device=None,
                          pad_mode: str = "reflect", **kwargs):
    """reference `mel_spec_auto_encoder` (:62-142).  wav: 1-D numpy array or tensor.  Returns a CUDA tensor (n_mels, T) in
    [0, 1], or with `cut` the list of (n_mels, partial_utterance_n_frames) views the reference returns."""
    w = _to_device(wav, device)
    mel_slices = None
    if cut:
        wave_slices, mel_slices = compute_partial_slices(
            w.shape[-1], sr=sr, mel_window_step=kwargs.pop("mel_window_step", AE_SPEC["mel_window_step"]),
            partial_utterance_n_frames=kwargs.pop("partial_utterance_n_frames", AE_SPEC["partial_utterance_n_frames"]), **kwargs)
        if wave_slices[-1].stop >= w.shape[-1]:
            w = _pad_to(w, wave_slices[-1].stop)
    fe = _frontend(w.device, sr=sr, n_fft=n_fft, hop_length=hop_length, win_length=window_length, n_mels=n_mels, fmin=fmin,
                   power=1, log_normalize=True, pad_mode=pad_mode, frames_major=False)
    mel = fe(w)
    if cut:
        return [mel[:, s] for s in mel_slices]
    return mel


def mel_spec_speaker_encoder(wav, sr: int = SE_SPEC["sr"], n_mels: int = SE_SPEC["n_mels"],
                             mel_window_length: float = SE_SPEC["mel_window_length"],
                             mel_window_step: float = SE_SPEC["mel_window_step"], cut: bool = False,
                             return_slices: bool = False, device=None, pad_mode: str = "reflect", **kwargs):
    """reference `mel_spec_speaker_encoder` (:144-217): power mel spectrogram, (T, n_mels); with `cut` the list of
    (partial_utterance_n_frames, n_mels) partial utterances `SpeakerEncoder.embed_utterance` batches (model.py:181)."""
    w = _to_device(wav, device)
    wave_slices = mel_slices = None
    if cut:
        wave_slices, mel_slices = compute_partial_slices(w.shape[-1], sr=sr, **kwargs)
        if wave_slices[-1].stop >= w.shape[-1]:
            w = _pad_to(w, wave_slices[-1].stop)
    fe = _frontend(w.device, sr=sr, n_fft=int(sr * mel_window_length / 1000), hop_length=int(sr * mel_window_step / 1000),
                   win_length=None, n_mels=n_mels, fmin=0.0, power=2, log_normalize=False, pad_mode=pad_mode, frames_major=True)
    mel = fe(w)
    if cut:
        mel = [mel[s] for s in mel_slices]
    if return_slices:
        return mel, wave_slices, mel_slices
    return mel


def mel_spectrogram(wav, model: str, **kwargs):
    """reference `mel_spectrogram` (:219-246)."""
    if model == "auto_encoder":
        return mel_spec_auto_encoder(wav, **kwargs)
    if model == "speaker_encoder":
        return mel_spec_speaker_encoder(wav, **kwargs)
    raise ValueError(f"'{model}' is not a valid model type. Must be either 'auto_encoder' or 'speaker_encoder'")
