"""Generate the mel front-end fixtures under `tests/golden/`.  Build-container only (needs /root/reference).

    python oracle/make_golden_mel.py

* `partial_slices.npz`   - outputs of the UNMODIFIED reference `compute_partial_slices`
                           (`/root/reference/autovc/audio/spectrogram.py:248-311`; pure numpy, imported with a librosa stub)
                           for a grid of waveform lengths and parameter sets, incl. both models' defaults.
* `mel_frontend_wav.npz` - a 1.5 s excerpt of the reference's own sample `data/samples/mette_183.wav` (22,050 Hz mono) plus a
                           16 kHz decimation of it, with the mel spectrograms the oracle restatement (`oracle/mel_oracle.py`)
                           computes for them.  librosa is absent, so these are REGRESSION vectors of the restatement, not
                           outputs of the reference - the restatement itself is pinned against torch.stft / torchaudio in
                           tests/test_mel_oracle.py.
"""
from __future__ import annotations

import os
import sys
import types

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
REF = "/root/reference"
OUT = os.path.join(ROOT, "tests", "golden")

SLICE_CASES = [  # (n_samples, sr, partial_utterance_n_frames, min_pad_coverage, overlap, mel_window_step)
    (16000, 16000, 160, 0.75, 0.5, 10), (40000, 16000, 160, 0.75, 0.5, 10), (47999, 16000, 160, 0.75, 0.5, 10),
    (48000, 16000, 160, 0.75, 0.5, 10), (100, 16000, 160, 0.75, 0.5, 10), (123457, 16000, 160, 0.5, 0.25, 10),
    (110000, 22050, 400, 0.75, 0.5, 12.5), (330750, 22050, 400, 0.75, 0.5, 12.5), (50000, 22050, 400, 0.75, 0.0, 12.5),
    (250000, 22050, 128, 1.0, 0.9, 12.5), (22050, 22050, 400, 0.75, 0.5, 12.5),
]


def main():
    sys.modules.setdefault("librosa", types.ModuleType("librosa"))
    sys.path.insert(0, REF)
    import importlib.util
    # the package __init__ pulls in audio I/O libraries that are absent: load the module file on its own
    for name in ("autovc", "autovc.utils"):
        sys.modules.setdefault(name, types.ModuleType(name))
    spec_h = importlib.util.spec_from_file_location("autovc.utils.hparams", os.path.join(REF, "autovc/utils/hparams.py"))
    hp = importlib.util.module_from_spec(spec_h); spec_h.loader.exec_module(hp)
    sys.modules["autovc.utils.hparams"] = hp
    spec_s = importlib.util.spec_from_file_location("ref_spectrogram", os.path.join(REF, "autovc/audio/spectrogram.py"))
    ref = importlib.util.module_from_spec(spec_s); spec_s.loader.exec_module(ref)

    rows = []
    for case in SLICE_CASES:
        n, sr, pun, cov, ov, step = case
        ws, ms = ref.compute_partial_slices(n, sr, partial_utterance_n_frames=pun, min_pad_coverage=cov, overlap=ov, mel_window_step=step)
        rows.append(dict(case=np.array(case, dtype=np.float64),
                         wav=np.array([[int(s.start), int(s.stop)] for s in ws], dtype=np.int64),
                         mel=np.array([[int(s.start), int(s.stop)] for s in ms], dtype=np.int64)))
    flat = {}
    for i, r in enumerate(rows):
        for k, v in r.items():
            flat[f"{k}_{i}"] = v
    np.savez_compressed(os.path.join(OUT, "partial_slices.npz"), n_cases=len(rows), **flat)
    # the reference's dB / normalisation helpers on a grid (pure numpy)
    grid = np.concatenate([[0.0, 1e-7, 1e-5], np.logspace(-5, 2, 50)]).astype(np.float32)
    db = ref.amp_to_db(grid)
    np.savez_compressed(os.path.join(OUT, "spec_helpers.npz"), grid=grid, db=db, norm=ref.normalize_spec(db),
                        denorm=ref.denormalize_spec(ref.normalize_spec(db)), amp=ref.db_to_amp(db))

    from scipy.io import wavfile
    from scipy.signal import resample_poly
    from oracle import mel_oracle as M
    sr, wav = wavfile.read(os.path.join(REF, "data/samples/mette_183.wav"))
    assert sr == 22050
    wav = wav.astype(np.float32) if wav.dtype.kind == "f" else wav.astype(np.float32) / 32768.0
    # stored as float16 (the fixture stays small); the vectors are computed from exactly the stored values
    w22 = wav[22050:22050 + 33075].astype(np.float16).astype(np.float32)          # 1.5 s from second 1
    w16 = resample_poly(w22.astype(np.float64), 320, 441)[:24000].astype(np.float16).astype(np.float32)
    ae = M.mel_spec_auto_encoder(w22)
    se = M.mel_spec_speaker_encoder(w16)
    ae_cut = np.stack(M.mel_spec_auto_encoder(w22, cut=True, partial_utterance_n_frames=64))
    se_cut = np.stack(M.mel_spec_speaker_encoder(w16, cut=True))
    np.savez_compressed(os.path.join(OUT, "mel_frontend_wav.npz"), w22=w22.astype(np.float16), w16=w16.astype(np.float16),
                        ae=ae.astype(np.float32), se=se.astype(np.float32), ae_cut=ae_cut.astype(np.float32),
                        se_cut=se_cut.astype(np.float32))
    print("AE mel", ae.shape, "SE mel", se.shape, "cuts", ae_cut.shape, se_cut.shape)


if __name__ == "__main__":
    main()
