"""Mel front-end kernels on the GPU against the CPU oracle (`oracle/mel_oracle.py`, librosa semantics) on the same inputs.

Tolerances: the transform runs in float64 on both sides and the rest in float32, so the differences are summation order
and log10f: normalised generator mel (range [0,1]) max-abs <= 2e-6 (measured 2e-7); speaker-encoder power mel <= 1e-5 relative to the
spectrogram's maximum (it is a linear-scale quantity spanning ~10 decades)."""
import numpy as np
import pytest
import torch

from conftest import load_golden
from oracle import mel_oracle as M

pytestmark = pytest.mark.gpu

AE_TOL = 2e-6
SE_RTOL = 1e-5


def _signal(n, seed, sr):
    rng = np.random.default_rng(seed)
    t = np.arange(n) / sr
    y = 0.3 * np.sin(2 * np.pi * 220 * t) + 0.1 * np.sin(2 * np.pi * 3300 * t + 1.0) + 0.01 * rng.normal(size=n)
    y[: n // 7] *= 1e-3                                   # a quiet stretch: exercises the low end of the dB scale
    return y.astype(np.float32)


def test_generator_mel_matches_oracle_golden_wav():
    from autovc_b200 import audio
    g = load_golden("mel_frontend_wav")
    w22 = g["w22"].astype(np.float32)
    mel = audio.mel_spec_auto_encoder(w22)
    assert mel.is_cuda and mel.shape == (80, 121)
    assert float((mel.cpu() - torch.from_numpy(g["ae"])).abs().max()) <= AE_TOL
    cut = audio.mel_spec_auto_encoder(w22, cut=True, partial_utterance_n_frames=64)
    assert len(cut) == 3 and all(c.shape == (80, 64) for c in cut)
    assert float((torch.stack(cut).cpu() - torch.from_numpy(g["ae_cut"])).abs().max()) <= AE_TOL


def test_speaker_encoder_mel_matches_oracle_golden_wav():
    from autovc_b200 import audio
    g = load_golden("mel_frontend_wav")
    w16 = g["w16"].astype(np.float32)
    mel = audio.mel_spec_speaker_encoder(w16)
    assert mel.shape == (151, 40)
    assert float((mel.cpu() - torch.from_numpy(g["se"])).abs().max()) <= SE_RTOL * float(g["se"].max())
    parts, ws, ms = audio.mel_spec_speaker_encoder(w16, cut=True, return_slices=True)
    assert len(parts) == 1 and parts[0].shape == (160, 40) and ws[0] == slice(0, 25600) and ms[0] == slice(0, 160)
    assert float((parts[0].cpu() - torch.from_numpy(g["se_cut"][0])).abs().max()) <= SE_RTOL * float(g["se_cut"].max())


@pytest.mark.parametrize("n,pad", [(22050, "reflect"), (1025, "reflect"), (35277, "reflect"), (30000, "constant"), (275 * 128 - 1, "reflect")])
def test_generator_mel_matches_live_oracle(n, pad):
    from autovc_b200 import audio
    y = _signal(n, n, 22050)
    ref = M.mel_spec_auto_encoder(y, pad_mode=pad)
    mel = audio.mel_spec_auto_encoder(torch.from_numpy(y).cuda(), pad_mode=pad)
    assert mel.shape == ref.shape == (80, 1 + n // 275)
    err = float((mel.cpu() - torch.from_numpy(ref)).abs().max())
    print(f"n={n} pad={pad}: max-abs {err:.2e}")
    assert err <= AE_TOL


@pytest.mark.parametrize("n", [16000, 201, 40001])
def test_speaker_encoder_mel_matches_live_oracle(n):
    from autovc_b200 import audio
    y = _signal(n, n, 16000)
    ref = M.mel_spec_speaker_encoder(y)
    mel = audio.mel_spec_speaker_encoder(y)
    assert mel.shape == ref.shape == (1 + n // 160, 40)
    assert float((mel.cpu() - torch.from_numpy(ref)).abs().max()) <= SE_RTOL * float(ref.max())


def test_batch_of_utterances_and_silence():
    """(B, n) input: every row equals the row computed alone; digital silence maps to 0 (the 1e-5 floor = -100 dB)."""
    from autovc_b200 import audio
    ys = np.stack([_signal(20000, s, 22050) for s in range(3)] + [np.zeros(20000, np.float32)])
    fe = audio.MelFrontend(22050, 2048, 275, 1100, 80, 40.0, power=1, log_normalize=True, device_index=0)
    both = fe(torch.from_numpy(ys).cuda())
    assert both.shape == (4, 80, 73)
    for i in range(4):
        assert torch.equal(both[i], fe(torch.from_numpy(ys[i]).cuda()))
    assert float(both[3].abs().max()) == 0.0
    assert float((both[0].cpu() - torch.from_numpy(M.mel_spec_auto_encoder(ys[0]))).abs().max()) <= AE_TOL
    with pytest.raises(ValueError):
        fe(torch.zeros(1000).cuda())                       # reflect padding needs > n_fft/2 samples, like np.pad / librosa


def test_wav_to_conversion_stays_on_the_device():
    """The wire format between the front-end and the generator is x[B,80,T] in [0,1]: wav -> mel -> chunks -> forward."""
    import autovc_b200
    from autovc_b200 import audio
    from autovc_b200.synth import synth_inputs, synth_state_dict
    ae = autovc_b200.AutoEncoder(device="cuda", verbose=False)
    ae.load_state_dict(synth_state_dict(0))
    y = _signal(22050 * 3, 5, 22050)
    chunks = audio.mel_spec_auto_encoder(y, cut=True, partial_utterance_n_frames=128)
    x = torch.stack(chunks)
    assert x.is_cuda and x.shape[1:] == (80, 128)
    _, c_org, c_trg = synth_inputs(1, 128, seed=2)
    out = ae.batch_forward(x, c_org.cuda(), c_trg.cuda(), overlap=0.5)
    assert out.shape == (80, 128 + (x.shape[0] - 1) * 64) and torch.isfinite(out).all()
