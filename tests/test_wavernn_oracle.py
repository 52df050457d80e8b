"""The WaveRNN oracle against golden vectors produced by the unmodified reference module (oracle/make_golden_wavernn.py)
and the host-side fold / cross-fade helpers of the product against the oracle.  CPU only."""
import os

import numpy as np
import pytest

from autovc_b200.synth import synth_uniforms, synth_wavernn_state_dict
from oracle import wavernn_oracle as wo

GOLD = os.path.join(os.path.dirname(__file__), "golden")
CASES = ["wavernn_t24_folded", "wavernn_t22_single", "wavernn_t22_exactfold"]


def _case(name):
    g = np.load(os.path.join(GOLD, name + ".npz"))
    sd = synth_wavernn_state_dict(int(g["wseed"]))
    mel = np.random.Generator(np.random.PCG64(int(g["mseed"]))).uniform(0, 1, size=(1, 80, int(g["T"]))).astype(np.float32)
    return g, sd, mel


@pytest.mark.parametrize("name", CASES)
def test_upsampling_matches_reference(name):
    g, sd, mel = _case(name)
    m, aux = wo.upsample(sd, np.pad(mel, ((0, 0), (0, 0), (2, 2))))
    assert m.shape == (1, int(g["T"]) * 275, 80) and aux.shape == (1, int(g["T"]) * 275, 128)
    np.testing.assert_allclose(m[0, ::37, ::7], g["up_m"], atol=2e-6)
    np.testing.assert_allclose(aux[0, ::275, ::5], g["up_aux"], atol=2e-5)


@pytest.mark.parametrize("name", CASES)
def test_generate_matches_reference(name):
    """Same weights, same random stream: the per-fold samples agree with the reference's for the whole run (float32 GEMV
    summation order differs between numpy and torch, so a mixture choice could flip on an exact tie; none does here)."""
    g, sd, mel = _case(name)
    batched, target, overlap = bool(g["batched"]), int(g["target"]), int(g["overlap"])
    mels, aux, wave_len = wo.condition(sd, mel, batched, target, overlap)
    B, L = g["samples"].shape
    assert mels.shape[:2] == (B, L)
    samples = wo.generate_rows(sd, mels, aux, synth_uniforms(L, B, int(g["useed"])))
    err = np.abs(samples - g["samples"])
    assert err.max() < 2e-3, (err.max(), int(np.argmax(err.max(axis=0))))
    wav = wo.finish(samples, wave_len, batched, target, overlap)
    assert wav.shape == g["wav"].shape and wav.dtype == np.float32
    np.testing.assert_allclose(wav, g["wav"], atol=2e-3)


def test_forced_run_reproduces_free_run():
    g, sd, mel = _case("wavernn_t24_folded")
    mels, aux, _ = wo.condition(sd, mel, True, int(g["target"]), int(g["overlap"]))
    B, L = g["samples"].shape
    U = synth_uniforms(L, B, int(g["useed"]))
    n = 200
    free = wo.generate_rows(sd, mels[:, :n], aux[:, :n], U[:n])
    forced, logits = wo.generate_rows(sd, mels[:, :n], aux[:, :n], U[:n], forced_x=free.T.copy(), return_logits=True)
    np.testing.assert_array_equal(free, forced)
    assert logits.shape == (n, B, 30) and np.isfinite(logits).all()


@pytest.mark.parametrize("total,target,overlap", [(6600, 1500, 100), (5500, 2600, 100), (35200, 11000, 550), (300, 1000, 50)])
def test_fold_and_unfold_shapes(total, target, overlap):
    x = np.random.default_rng(0).standard_normal((1, total, 3)).astype(np.float32)
    f = wo.fold_with_overlap(x, target, overlap)
    nf, padding = wo.fold_params(total, target, overlap)
    assert f.shape == (nf, target + 2 * overlap, 3)
    for i in range(nf):
        s = i * (target + overlap)
        e = min(s + target + 2 * overlap, total)
        np.testing.assert_array_equal(f[i, :e - s], x[0, s:e])
        assert not f[i, e - s:].any()
    y = np.ones((nf, target + 2 * overlap))
    u = wo.xfade_and_unfold(y, target, overlap)
    assert u.shape == (nf * (target + overlap) + overlap,)
    # equal-power cross-fade: fade_in^2 + fade_out^2 == 1 over the faded part of every seam
    if nf > 1:
        seam = u[target + overlap + overlap // 2: target + 2 * overlap]
        t = np.linspace(-1, 1, overlap - overlap // 2)
        np.testing.assert_allclose(seam, np.sqrt(0.5 * (1 + t)) + np.sqrt(0.5 * (1 - t)), atol=1e-12)
