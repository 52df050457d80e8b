"""d-vector speaker encoder on the GPU against the golden vectors generated from the unmodified
reference `SpeakerEncoder` (oracle/make_golden.py) and against the CPU oracle on other seeded inputs.

Tolerances: fp32 validation mode max-abs <= 1e-4 on the unit-norm embeddings; bf16 mode max-abs <= 1e-3 (measured ~3e-4) and
cosine >= 0.999 per row (an embedding row has norm 1, so 1e-3 max-abs is ~2e-2 relative to the typical |element| of
1/16 - the same bf16 operand-rounding budget as the generator's content codes)."""
import numpy as np
import pytest
import torch

from autovc_b200.synth import synth_speaker_encoder
from conftest import load_golden

pytestmark = pytest.mark.gpu

FP32_TOL = 1e-4
BF16_TOL = 1e-3
BF16_COS = 0.999

_SE = {}


def encoder(precision, seed=5):
    import autovc_b200
    key = (precision, seed)
    if key not in _SE:
        m = autovc_b200.SpeakerEncoder(device="cuda", verbose=False, precision=precision)
        m.load_state_dict(synth_speaker_encoder(seed)[0])
        _SE[key] = m
    return _SE[key]


def _check(got, ref, precision):
    err = float((got.cpu() - ref).abs().max())
    cos = float(torch.nn.functional.cosine_similarity(got.cpu(), ref, dim=-1).min())
    print(f"{precision}: max-abs {err:.2e} cos {cos:.6f}")
    if precision == "fp32":
        assert err <= FP32_TOL
    else:
        assert err <= BF16_TOL and cos >= BF16_COS


@pytest.mark.parametrize("precision", ["fp32", "bf16"])
def test_forward_matches_reference_golden(precision):
    g = load_golden("speaker_encoder_p8")
    _, parts = synth_speaker_encoder(int(g["seed"]))
    m = encoder(precision)
    emb = m(parts.cuda())
    assert emb.shape == (8, 256)
    assert float((emb.norm(dim=1) - 1).abs().max()) <= 1e-5
    _check(emb, torch.from_numpy(g["partial_embeds"]), precision)
    utt, partials = m.embed_partials(parts, return_partials=True)
    assert utt.shape == (256,) and torch.equal(partials, emb.cpu())
    _check(utt, torch.from_numpy(g["embed"]), precision)
    assert torch.equal(m.embed_frames_batch(parts.numpy()), emb.cpu())


@pytest.mark.parametrize("precision", ["fp32", "bf16"])
@pytest.mark.parametrize("N,T", [(1, 160), (3, 1), (5, 37), (70, 160)])
def test_forward_matches_live_oracle(precision, N, T):
    """Ragged shapes: one partial, one frame, T off the partial length, more partials than one 64-row batch tile."""
    from oracle import autovc_oracle as O
    sd, _ = synth_speaker_encoder(11)
    parts = torch.from_numpy(np.random.default_rng(100 + N + T).normal(size=(N, T, 40)).astype(np.float32))
    ref = O.speaker_encoder_forward(sd, parts)
    _check(encoder(precision, 11)(parts.cuda()), ref, precision)


@pytest.mark.parametrize("precision", ["fp32", "bf16"])
def test_embed_speakers_batches_utterances(precision):
    """Utterances with different partial counts in one batch equal the same utterances embedded one at a time
    (rows of a batch are independent) and the oracle's embed_utterance."""
    from oracle import autovc_oracle as O
    sd, _ = synth_speaker_encoder(5)
    rng = np.random.default_rng(77)
    groups = [torch.from_numpy(rng.normal(size=(n, 160, 40)).astype(np.float32)) for n in (3, 1, 6)]
    m = encoder(precision)
    both = m.embed_speakers(groups)
    assert both.shape == (3, 256)
    for i, grp in enumerate(groups):
        alone = m.embed_partials(grp)
        assert float((both[i].cpu() - alone).abs().max()) <= 1e-5
        _check(both[i], O.embed_utterance_from_partials(sd, grp), precision)
    mean = m.learn_speaker("spk", groups)
    assert torch.allclose(mean, both.mean(dim=0).cpu(), atol=1e-6) and "spk" in m.speakers


def test_feeds_voice_converter():
    """The encoder is the `embed_speaker` hook of VoiceConverter: c_org / c_trg rows come from the GPU d-vector path."""
    import autovc_b200
    from autovc_b200.synth import synth_inputs, synth_state_dict
    ae = autovc_b200.AutoEncoder(device="cuda", verbose=False, precision="bf16")
    ae.load_state_dict(synth_state_dict(0))
    se = encoder("bf16")
    rng = np.random.default_rng(3)
    spk = {k: rng.normal(size=(4, 160, 40)).astype(np.float32) for k in ("a", "b")}
    vc = autovc_b200.VoiceConverter(ae, embed_speaker=se.as_embed_fn(), device="cuda")
    x, _, _ = synth_inputs(2, 128, seed=9)
    ca = vc.embedding("a", spk["a"])
    cb = vc.embedding("b", spk["b"])
    assert vc.embed_calls == 2 and ca.shape == (256,)
    out = vc.convert_batch([x[0], x[1]], [ca, cb], [cb, ca])
    _, post, _ = ae(x.cuda(), torch.stack([ca, cb]).cuda(), torch.stack([cb, ca]).cuda())
    for i in range(2):
        assert torch.equal(out[i].cpu(), post[i].cpu())


def test_errors():
    m = encoder("bf16")
    with pytest.raises(RuntimeError):
        m(torch.zeros(2, 160, 40))                  # CPU tensor: no fallback
    with pytest.raises(RuntimeError):
        m(torch.zeros(2, 160, 41).cuda())
    with pytest.raises(NotImplementedError):
        m(torch.zeros(2, 160, 40).cuda(), hidden_init=(None, None))
    import autovc_b200
    blank = autovc_b200.Engine((32, 256, 512, 32), 0, "bf16")
    with pytest.raises(RuntimeError):
        blank.se_forward(torch.zeros(1, 4, 40).cuda())   # weights never loaded
    with pytest.raises(KeyError):
        blank.load_state_dict({"lstm.weight_ih_l0": torch.zeros(1024, 40)}, speaker_encoder=True)
    blank.close()
