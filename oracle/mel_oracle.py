"""CPU oracle of the mel front-end - TEST INFRASTRUCTURE, never on the product path.

Restates, in numpy, what the reference computes in `autovc/audio/spectrogram.py`:

    mel_spec_auto_encoder   (:62-142)   stft(2048, hop 275, win 1100) -> |.| -> slaney mel (80, fmin 40) -> dB -> normalise
    mel_spec_speaker_encoder(:144-217)  melspectrogram(n_fft 400, hop 160, 40 mels, power 2), transposed
    compute_partial_slices  (:248-311)  integer arithmetic on sample / frame counts
    amp_to_db / normalize_spec / denormalize_spec (:14-60)

The arithmetic of the first two lives in a third-party dependency that is ABSENT here: librosa (`requirements.txt:7`,
unpinned; the pinned `numba==0.49.1` on line 15 dates the environment to librosa 0.8.x).  The functions below restate the
published algorithms of that version:

    librosa.stft              core/spectrum.py   centre-padded (pad_mode="reflect", the 0.8 default; "constant" from 0.10 on
                                                  is selectable), periodic Hann of win_length zero-padded to n_fft, rfft per
                                                  frame computed in float64 and stored as complex64
    librosa.filters.mel       filters.py         Slaney mel scale (htk=False), triangular filters, norm="slaney", float32
    librosa.feature.melspectrogram               mel_basis . |stft|**power

Pinning: librosa itself cannot be imported, so parity with it is UNPINNED; the restatement is pinned against two
independent implementations available in this image (tests/test_mel_oracle.py): `torch.stft` for the transform and
`torchaudio.functional.melscale_fbanks(norm="slaney", mel_scale="slaney")` for the filterbank, and
`compute_partial_slices` against golden vectors generated from the unmodified reference function
(`oracle/make_golden.py`, tests/golden/partial_slices.npz).
"""
from __future__ import annotations

import numpy as np

AE = dict(sr=22050, n_mels=80, n_fft=2048, hop_length=275, window_length=1100, fmin=40.0,
          mel_window_step=12.5, partial_utterance_n_frames=400)                  # reference utils/hparams.py:6-15
SE = dict(sr=16000, n_mels=40, mel_window_step=10, mel_window_length=25, partial_utterance_n_frames=160)   # :52-58


# ---- librosa.filters.mel (Slaney) -------------------------------------------------------------------------------------
def hz_to_mel(f):
    f = np.asanyarray(f, dtype=np.float64)
    f_sp = 200.0 / 3
    mels = f / f_sp
    min_log_hz, min_log_mel, logstep = 1000.0, 1000.0 / f_sp, np.log(6.4) / 27.0
    return np.where(f >= min_log_hz, min_log_mel + np.log(np.maximum(f, 1e-300) / min_log_hz) / logstep, mels)


def mel_to_hz(m):
    m = np.asanyarray(m, dtype=np.float64)
    f_sp = 200.0 / 3
    min_log_hz, min_log_mel, logstep = 1000.0, 1000.0 / f_sp, np.log(6.4) / 27.0
    return np.where(m >= min_log_mel, min_log_hz * np.exp(logstep * (m - min_log_mel)), f_sp * m)


def mel_filterbank(sr, n_fft, n_mels, fmin=0.0, fmax=None) -> np.ndarray:
    """(n_mels, 1 + n_fft//2) float32, as `librosa.filters.mel(sr, n_fft, n_mels, fmin, fmax, htk=False, norm='slaney')`."""
    fmax = sr / 2.0 if fmax is None else fmax
    fftfreqs = np.linspace(0, sr / 2.0, 1 + n_fft // 2)
    mel_f = mel_to_hz(np.linspace(hz_to_mel(fmin), hz_to_mel(fmax), n_mels + 2))
    fdiff = np.diff(mel_f)
    ramps = np.subtract.outer(mel_f, fftfreqs)
    weights = np.zeros((n_mels, 1 + n_fft // 2), dtype=np.float32)
    for i in range(n_mels):
        lower = -ramps[i] / fdiff[i]
        upper = ramps[i + 2] / fdiff[i + 1]
        weights[i] = np.maximum(0, np.minimum(lower, upper))
    enorm = 2.0 / (mel_f[2:n_mels + 2] - mel_f[:n_mels])
    weights *= enorm[:, np.newaxis]
    return weights


# ---- librosa.stft -------------------------------------------------------------------------------------------------------
def hann_periodic(n) -> np.ndarray:
    return 0.5 - 0.5 * np.cos(2.0 * np.pi * np.arange(n) / n)


def stft(y, n_fft, hop_length, win_length=None, pad_mode="reflect") -> np.ndarray:
    """(1 + n_fft//2, 1 + len(y)//hop) complex64."""
    y = np.asarray(y, dtype=np.float32)
    win_length = n_fft if win_length is None else win_length
    win = np.zeros(n_fft, dtype=np.float64)
    off = (n_fft - win_length) // 2
    win[off:off + win_length] = hann_periodic(win_length)
    # librosa: scipy get_window (float64) -> pad_center; the float64 window times the float32 frames is a float64 product
    if pad_mode == "reflect" and len(y) <= n_fft // 2:
        raise ValueError("reflect padding needs more than n_fft//2 samples")
    yp = np.pad(y, n_fft // 2, mode=pad_mode)
    n_frames = 1 + (len(yp) - n_fft) // hop_length
    idx = np.arange(n_fft)[:, None] + hop_length * np.arange(n_frames)[None, :]
    frames = yp[idx]                                  # (n_fft, n_frames) float32
    return np.fft.rfft(win[:, None] * frames, axis=0).astype(np.complex64)


def melspectrogram(y=None, sr=22050, S=None, n_fft=2048, hop_length=512, win_length=None, n_mels=128, fmin=0.0, fmax=None,
                   power=2.0, pad_mode="reflect") -> np.ndarray:
    if S is None:
        S = np.abs(stft(y, n_fft, hop_length, win_length, pad_mode)) ** power
    return np.dot(mel_filterbank(sr, n_fft, n_mels, fmin, fmax), S)


# ---- reference autovc/audio/spectrogram.py ----------------------------------------------------------------------------------
def amp_to_db(amplitude):                                    # :53-55
    return 20 * np.log10(np.maximum(1e-5, amplitude))


def normalize_spec(spectrogram, min_level_db=-100):          # :14-32
    return np.clip((spectrogram - min_level_db) / -min_level_db, 0, 1)


def denormalize_spec(spectrogram, min_level_db=-100):        # :34-51
    return (np.clip(spectrogram, 0, 1) * -min_level_db) + min_level_db


def compute_partial_slices(n_samples, sr, partial_utterance_n_frames=160, min_pad_coverage=0.75, overlap=0.5,
                           mel_window_step=10):              # :248-311
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


def mel_spec_auto_encoder(wav, sr=AE["sr"], cut=False, n_mels=AE["n_mels"], n_fft=AE["n_fft"], hop_length=AE["hop_length"],
                          window_length=AE["window_length"], fmin=AE["fmin"], pad_mode="reflect", **kwargs):   # :62-142
    wav = np.asarray(wav, dtype=np.float32)
    if cut:
        wave_slices, mel_slices = compute_partial_slices(
            len(wav), sr=sr, mel_window_step=kwargs.pop("mel_window_step", AE["mel_window_step"]),
            partial_utterance_n_frames=kwargs.pop("partial_utterance_n_frames", AE["partial_utterance_n_frames"]), **kwargs)
        max_wave_length = wave_slices[-1].stop
        if max_wave_length >= len(wav):
            wav = np.pad(wav, (0, max_wave_length - len(wav)), "constant")
    spectrogram = stft(wav, n_fft, hop_length, window_length, pad_mode)
    mel = melspectrogram(S=np.abs(spectrogram), sr=sr, n_fft=n_fft, n_mels=n_mels, fmin=fmin)
    mel = normalize_spec(amp_to_db(mel))
    if cut:
        return [mel[:, s] for s in mel_slices]
    return mel


def mel_spec_speaker_encoder(wav, sr=SE["sr"], n_mels=SE["n_mels"], mel_window_length=SE["mel_window_length"],
                             mel_window_step=SE["mel_window_step"], cut=False, return_slices=False, pad_mode="reflect",
                             **kwargs):                      # :144-217
    wav = np.asarray(wav, dtype=np.float32)
    wave_slices = mel_slices = None
    if cut:
        wave_slices, mel_slices = compute_partial_slices(len(wav), sr=sr, **kwargs)
        max_wave_length = wave_slices[-1].stop
        if max_wave_length >= len(wav):
            wav = np.pad(wav, (0, max_wave_length - len(wav)), "constant")
    mel = melspectrogram(wav, sr, n_fft=int(sr * mel_window_length / 1000), hop_length=int(sr * mel_window_step / 1000),
                         n_mels=n_mels, pad_mode=pad_mode)
    mel = mel.astype(np.float32).T
    if cut:
        mel = [mel[s] for s in mel_slices]
    if return_slices:
        return mel, wave_slices, mel_slices
    return mel
