"""WaveRNN vocoder: the CUDA path through the C ABI against the numpy oracle and against golden
vectors produced by the unmodified reference module.

Tolerances.  The device path computes with bf16 operands (weights and the exchanged activations) and fp32 accumulation /
state; the oracle is float32 throughout.  Element-wise bounds are stated on a TEACHER-FORCED run (the oracle's samples fed
back), where they mean something: |logits| error <= 2e-2 and |sample| error <= 2e-2 (measured: 3e-3 / 2e-3) wherever the
mixture choice is not a near-tie.  A free run
feeds its own samples back through 2 GRUs for thousands of steps; its error is bounded statistically (median, tail),
and exactly for the first steps.  The conditioning network is fp32: <= 1e-5.
"""
import ctypes
import os

import numpy as np
import pytest
import torch

from autovc_b200 import _lib as L
from autovc_b200.synth import synth_uniforms, synth_wavernn_state_dict
from autovc_b200.vocoder import WaveRNN, fold_params
from oracle import wavernn_oracle as wo

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(__file__), "golden")


@pytest.fixture(scope="module")
def sd():
    return synth_wavernn_state_dict(11)


@pytest.fixture(scope="module")
def voc(sd):
    m = WaveRNN(device="cuda")
    m.load_state_dict(sd)
    return m


def _mel(T, seed=3, U=1):
    return np.random.Generator(np.random.PCG64(seed)).uniform(0, 1, size=(U, 80, T)).astype(np.float32)


@pytest.mark.parametrize("U,T", [(1, 24), (3, 37), (2, 16)])
def test_conditioning_matches_oracle(voc, sd, U, T):
    mel = _mel(T, seed=5, U=U)
    m_ref, aux_ref = wo.upsample(sd, np.pad(mel, ((0, 0), (0, 0), (2, 2))))
    m_up, aux = voc.condition(torch.from_numpy(mel).cuda())
    assert m_up.shape == (U, T * 275, 80) and aux.shape == (U, T, 128)
    np.testing.assert_allclose(m_up.cpu().numpy(), m_ref, atol=1e-5)
    np.testing.assert_allclose(aux.cpu().numpy(), aux_ref[:, ::275], atol=1e-5)
    # Stretch2d(275): the reference's aux is constant over each frame
    assert np.array_equal(aux_ref[:, 274], aux_ref[:, 0]) and np.array_equal(aux_ref[:, 275], aux_ref[:, 549])


def _folded(sd, mel, target, overlap):
    mels_f, aux_f, wave_len = wo.condition(sd, mel, True, target, overlap)
    B = mels_f.shape[0]
    return mels_f, aux_f, [i * (target + overlap) for i in range(B)]


def test_forced_run_matches_oracle(voc, sd):
    """Teacher forcing: every step's logits and sample against the oracle, including the padded tail of the last fold."""
    mel = _mel(24)
    target, overlap, n = 1500, 100, 400
    mels_f, aux_f, starts = _folded(sd, mel, target, overlap)
    B = len(starts)
    # the last fold runs off the end of the conditioning after 200 steps (zeros: pad_tensor 'after')
    off = 0
    U = synth_uniforms(n, B, 21)
    ref, lg_ref = wo.generate_rows(sd, mels_f[:, off:off + n], aux_f[:, off:off + n], U, return_logits=True)
    assert not mels_f[-1, off + n - 1].any() and mels_f[-1, off].any()
    m_up, aux = voc.condition(torch.from_numpy(mel).cuda())
    s, lg = voc.sample_rows(m_up, aux, [0] * B, [st + off for st in starts], n, uniforms=U, forced_x=ref.T.copy(), return_logits=True)
    assert np.abs(lg.cpu().numpy() - lg_ref).max() <= 2e-2
    # the sample is continuous in the logits except for the mixture choice (an argmax over Gumbel-perturbed logits): where
    # the oracle's two best candidates are closer than twice the logits bound the choice may legitimately flip
    pert = lg_ref[:, :, :10] - np.log(-np.log(U[:, :, :10]))
    top2 = np.sort(pert, axis=2)[:, :, -2:]
    near_tie = ((top2[:, :, 1] - top2[:, :, 0]) < 4e-2).T            # (rows, steps)
    err = np.abs(s.cpu().numpy() - ref)
    assert near_tie.mean() < 0.06           # a property of the oracle and the random stream alone (measured 0.039)
    assert err[~near_tie].max() <= 2e-2


@pytest.mark.parametrize("name", ["wavernn_t24_folded", "wavernn_t22_exactfold", "wavernn_t22_single"])
def test_generate_against_reference_golden(voc, name):
    """`generate` end to end (conditioning, fold windows, loop, cross-fade, fade-out) against the waveform the unmodified
    reference produced from the same weights and random stream."""
    g = np.load(os.path.join(GOLD, name + ".npz"))
    T, target, overlap, batched = int(g["T"]), int(g["target"]), int(g["overlap"]), bool(g["batched"])
    mel = _mel(T, seed=int(g["mseed"]))
    B, Lsteps = g["samples"].shape
    U = synth_uniforms(Lsteps, B, int(g["useed"]))
    wav = voc.generate(torch.from_numpy(mel), batched=batched, target=target, overlap=overlap, uniforms=U)
    assert wav.dtype == np.float32 and wav.shape == g["wav"].shape
    assert voc.last_rows == B and voc.last_steps == Lsteps
    err = np.abs(wav - g["wav"])
    n0 = 0 if not batched else overlap        # the first overlap//2 samples of a folded waveform are silenced by the cross-fade
    assert err[n0:n0 + 20].max() <= 1e-2
    assert np.median(err) <= 5e-3, np.median(err)
    assert np.quantile(err, 0.99) <= 0.15, np.quantile(err, 0.99)
    assert np.corrcoef(wav, g["wav"])[0, 1] >= 0.98


def test_rows_do_not_depend_on_the_batch(voc, sd):
    """A row's samples are bitwise the same alone, in a batch of 5, among 200 rows (2 tiles) and among 300 (4 tiles)."""
    mel = _mel(30, seed=9)
    m_up, aux = voc.condition(torch.from_numpy(mel).cuda())
    n = 120
    rng = np.random.default_rng(0)
    starts = rng.integers(0, 30 * 275 - 50, size=300).astype(np.int32)
    U = synth_uniforms(n, 300, 4)
    full = voc.sample_rows(m_up, aux, [0] * 300, starts, n, uniforms=U).cpu().numpy()
    for rows in (1, 5, 200):
        part = voc.sample_rows(m_up, aux, [0] * rows, starts[:rows], n, uniforms=U[:, :rows].copy()).cpu().numpy()
        assert np.array_equal(part, full[:rows]), rows
    assert np.isfinite(full).all() and np.abs(full).max() <= 1.0


def test_more_rows_than_one_launch(voc):
    """Beyond avc_voc_max_rows() the work list is split into launches; results do not depend on the split."""
    maxr = int(L.lib().avc_voc_max_rows())
    assert maxr == 512
    mel = _mel(21, seed=2, U=2)
    m_up, aux = voc.condition(torch.from_numpy(mel).cuda())
    rows, n = maxr + 40, 60
    utt = (np.arange(rows) % 2).astype(np.int32)
    starts = ((np.arange(rows) * 53) % (21 * 275)).astype(np.int32)
    U = synth_uniforms(n, rows, 8)
    full = voc.sample_rows(m_up, aux, utt, starts, n, uniforms=U).cpu().numpy()
    assert voc.last_launch_count == 6
    tail = voc.sample_rows(m_up, aux, utt[maxr:], starts[maxr:], n, uniforms=U[:, maxr:].copy()).cpu().numpy()
    assert np.array_equal(full[maxr:], tail)
    # the two utterances give different audio
    assert not np.array_equal(full[0], full[1])


def test_internal_random_stream(voc):
    mel = _mel(21, seed=4)
    m_up, aux = voc.condition(torch.from_numpy(mel).cuda())
    a = voc.sample_rows(m_up, aux, [0, 0, 0], [0, 0, 500], 300, seed=1).cpu().numpy()
    b = voc.sample_rows(m_up, aux, [0, 0, 0], [0, 0, 500], 300, seed=1).cpu().numpy()
    c = voc.sample_rows(m_up, aux, [0, 0, 0], [0, 0, 500], 300, seed=2).cpu().numpy()
    assert np.array_equal(a, b) and not np.array_equal(a, c)
    assert not np.array_equal(a[0], a[1])          # same conditioning, different rows: independent draws
    assert np.abs(a).max() <= 1.0 and 0.05 < a.std() < 0.6


def test_generate_batch_equals_single_generates(voc):
    mels = _mel(23, seed=6, U=3)
    target, overlap = 2000, 150
    nf, _ = fold_params(23 * 275, target, overlap)
    U = synth_uniforms(target + 2 * overlap, 3 * nf, 12)
    waves = voc.generate_batch([m for m in mels], True, target, overlap, uniforms=U)
    assert len(waves) == 3 and all(w.shape == (22 * 275,) for w in waves)
    one = voc.generate(mels[1:2], True, target, overlap, uniforms=U[:, nf:2 * nf].copy())
    assert np.array_equal(one, waves[1])
    # the reference's 20-frame fade-out ends in silence
    assert waves[0][-1] == 0.0


def test_too_short_mel_and_bad_arguments(voc):
    with pytest.raises(ValueError):
        voc.generate(_mel(20))                     # the reference's fade-out fails on < 21 frames (model.py:327)
    with pytest.raises(RuntimeError):
        voc.condition(torch.zeros(1, 80, 30))      # CPU tensor: no fallback
    h = ctypes.c_void_p()
    assert L.lib().avc_voc_create(ctypes.byref(h), 0) == L.AVC_OK
    dummy = torch.zeros(16, device="cuda")
    rc = L.lib().avc_voc_generate(h, dummy.data_ptr(), dummy.data_ptr(), 1, 1, dummy.data_ptr(), dummy.data_ptr(), 1, 1, None, 0, None,
                                  dummy.data_ptr(), None, dummy.data_ptr(), None)
    assert rc == L.AVC_ERR_NOT_LOADED and b"load_weights" in L.lib().avc_voc_last_error(h)
    L.lib().avc_voc_destroy(h)


def test_missing_weight_is_reported(sd):
    h = ctypes.c_void_p()
    assert L.lib().avc_voc_create(ctypes.byref(h), 0) == L.AVC_OK
    keep = {k: v.float().contiguous() for k, v in sd.items() if k not in ("step", "fc2.bias") and not k.endswith("num_batches_tracked")}
    table = (L.AvcWeightDesc * len(keep))()
    for i, (k, v) in enumerate(keep.items()):
        table[i] = L.AvcWeightDesc(k.encode(), ctypes.cast(v.data_ptr(), ctypes.POINTER(ctypes.c_float)), v.numel())
    assert L.lib().avc_voc_load_weights(h, table, len(keep)) == L.AVC_ERR_MISSING_WEIGHT
    assert b"fc2.bias" in L.lib().avc_voc_last_error(h)
    L.lib().avc_voc_destroy(h)
