"""CPU tests of the mel front-end's oracle and host logic (no GPU).

librosa is absent from this image, so the numpy restatement in `oracle/mel_oracle.py` is pinned against two independent
implementations (torch.stft, torchaudio's Slaney filterbank), the integer slicing logic against golden vectors generated from
the unmodified reference function, and the library's host-built filterbank against the oracle bit for bit."""
import numpy as np
import pytest
import torch

from conftest import load_golden
from oracle import mel_oracle as M


def _noise(n, seed=0, scale=0.1):
    return (np.random.default_rng(seed).normal(size=n) * scale).astype(np.float32)


@pytest.mark.parametrize("n_fft,hop,win,pad", [(2048, 275, 1100, "reflect"), (400, 160, 400, "reflect"), (2048, 275, 1100, "constant")])
def test_stft_matches_torch(n_fft, hop, win, pad):
    y = _noise(22050 + 77)
    S = M.stft(y, n_fft, hop, win, pad_mode=pad)
    w = torch.zeros(n_fft, dtype=torch.float64)
    off = (n_fft - win) // 2
    w[off:off + win] = torch.from_numpy(M.hann_periodic(win))
    T = torch.stft(torch.from_numpy(y).double(), n_fft, hop, n_fft, window=w, center=True, pad_mode=pad, return_complex=True).numpy()
    assert S.shape == T.shape == (1 + n_fft // 2, 1 + len(y) // hop)
    assert np.abs(S - T).max() <= 1e-6 * np.abs(T).max()


@pytest.mark.parametrize("sr,n_fft,n_mels,fmin", [(22050, 2048, 80, 40.0), (16000, 400, 40, 0.0)])
def test_filterbank_matches_torchaudio(sr, n_fft, n_mels, fmin):
    import torchaudio
    fb = M.mel_filterbank(sr, n_fft, n_mels, fmin)
    ta = torchaudio.functional.melscale_fbanks(1 + n_fft // 2, fmin, sr / 2, n_mels, sr, norm="slaney", mel_scale="slaney").numpy().T
    assert fb.dtype == np.float32 and fb.shape == ta.shape
    assert np.abs(fb - ta).max() <= 2e-7
    assert (fb.sum(axis=1) > 0).all()          # no empty filter at these resolutions


def test_partial_slices_match_reference_golden():
    g = load_golden("partial_slices")
    from autovc_b200.audio import compute_partial_slices as host_slices
    for i in range(int(g["n_cases"])):
        n, sr, pun, cov, ov, step = g[f"case_{i}"]
        for fn in (M.compute_partial_slices, host_slices):
            ws, ms = fn(int(n), int(sr), partial_utterance_n_frames=int(pun), min_pad_coverage=float(cov), overlap=float(ov),
                        mel_window_step=float(step))
            assert [[s.start, s.stop] for s in ws] == g[f"wav_{i}"].tolist()
            assert [[s.start, s.stop] for s in ms] == g[f"mel_{i}"].tolist()


def test_db_and_normalisation_helpers_match_reference_golden():
    g = load_golden("spec_helpers")
    from autovc_b200 import audio
    grid = g["grid"]
    assert np.allclose(M.amp_to_db(grid), g["db"], atol=1e-5)
    assert np.allclose(M.normalize_spec(g["db"]), g["norm"], atol=1e-7)
    t = torch.from_numpy(grid)
    assert np.allclose(audio.amp_to_db(t).numpy(), g["db"], atol=1e-4)
    assert np.allclose(audio.normalize_spec(torch.from_numpy(g["db"])).numpy(), g["norm"], atol=1e-6)
    assert np.allclose(audio.denormalize_spec(torch.from_numpy(g["norm"])).numpy(), g["denorm"], atol=1e-4)
    assert np.allclose(audio.db_to_amp(torch.from_numpy(g["db"])).numpy(), g["amp"], rtol=1e-4)


def test_oracle_regression_vectors():
    g = load_golden("mel_frontend_wav")
    w22, w16 = g["w22"].astype(np.float32), g["w16"].astype(np.float32)
    ae = M.mel_spec_auto_encoder(w22)
    assert ae.shape == (80, 121) and ae.min() >= 0 and ae.max() <= 1
    assert np.abs(ae - g["ae"]).max() <= 1e-6
    se = M.mel_spec_speaker_encoder(w16)
    assert se.shape == (151, 40)
    assert np.abs(se - g["se"]).max() <= 1e-6 * g["se"].max()
    cut = M.mel_spec_auto_encoder(w22, cut=True, partial_utterance_n_frames=64)
    assert np.abs(np.stack(cut) - g["ae_cut"]).max() <= 1e-6


def test_library_filterbank_is_the_oracles(built_lib):
    """Host-only handles (device = -1): the filterbank the kernels use is bit-identical to the oracle's."""
    import autovc_b200
    for args, ref in [((22050, 2048, 275, 1100, 80, 40.0), M.mel_filterbank(22050, 2048, 80, 40.0)),
                      ((16000, 400, 160, None, 40, 0.0), M.mel_filterbank(16000, 400, 40))]:
        fe = autovc_b200.MelFrontend(*args, device_index=-1)
        assert np.array_equal(fe.filterbank().numpy(), ref)
        assert fe.num_frames(22050) == 1 + 22050 // args[2]
        with pytest.raises((RuntimeError, ValueError)):
            fe(torch.zeros(4096))                          # CPU tensor: no fallback
        fe.close()
    with pytest.raises(RuntimeError):
        autovc_b200.MelFrontend(22050, 2048, 275, 4096, 80, device_index=-1)     # window longer than n_fft
