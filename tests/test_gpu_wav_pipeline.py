"""Waveform-level conversion on the GPU (`WavVoiceConverter`: mel front-end -> d-vector encoder -> generator -> vocoder,
reference voice_converter.py:140-175) against the chain of CPU oracles on the same waveforms and seeded weights.

Tolerances: the chain's error is the generator's (fp32 mode <= 1e-4, bf16 mode max-abs(mel_outputs_postnet) <= 1e-2) - the mel
front-end contributes 2e-7 and the embeddings 1e-4 / 3e-4 (fp32 / bf16), which the generator sees as inputs."""
import numpy as np
import pytest
import torch

from autovc_b200.synth import synth_speaker_encoder, synth_state_dict, synth_uniforms, synth_wavernn_state_dict

pytestmark = pytest.mark.gpu


def _signal(n, seed, sr=22050):
    rng = np.random.default_rng(seed)
    t = np.arange(n) / sr
    f0 = 110 + 40 * seed
    y = 0.3 * np.sin(2 * np.pi * f0 * t) + 0.1 * np.sin(2 * np.pi * (2900 + 100 * seed) * t + 1.0) + 0.02 * rng.normal(size=n)
    return (y * (0.5 + 0.5 * np.sin(2 * np.pi * 1.5 * t) ** 2)).astype(np.float32)


def _oracle_embed(se_sd, wav):
    from oracle import autovc_oracle as O, mel_oracle as M
    parts = np.stack(M.mel_spec_speaker_encoder(wav, cut=True))
    return O.embed_utterance_from_partials(se_sd, torch.from_numpy(parts))


def _oracle_convert(sd, se_sd, src, c_src_wav, trg):
    from oracle import autovc_oracle as O, mel_oracle as M
    x = torch.from_numpy(M.mel_spec_auto_encoder(src))[None]
    c_s, c_t = _oracle_embed(se_sd, c_src_wav)[None], _oracle_embed(se_sd, trg)[None]
    return O.generator_forward(sd, x, c_s, c_t)[1][0]


_CACHE = {}


def _stack(precision):
    import autovc_b200
    if precision not in _CACHE:
        ae = autovc_b200.AutoEncoder(device="cuda", verbose=False, precision=precision)
        ae.load_state_dict(synth_state_dict(0))
        se = autovc_b200.SpeakerEncoder(device="cuda", verbose=False, precision=precision)
        se.load_state_dict(synth_speaker_encoder(5)[0])
        _CACHE[precision] = (ae, se)
    return _CACHE[precision]


@pytest.mark.parametrize("precision", ["fp32", "bf16"])
def test_convert_multiple_matches_oracle_chain(precision):
    from autovc_b200.voice_converter import WavVoiceConverter
    ae, se = _stack(precision)
    vc = WavVoiceConverter(ae, se, device="cuda")
    sources = [_signal(48510, 1), _signal(30000, 2), _signal(48510, 3)]        # two length buckets
    targets = [_signal(40000, 4), _signal(52000, 5)]
    posts = vc.convert_multiple(sources, targets, match_method="all_combinations", vocode=False)
    assert len(posts) == 6
    sd, se_sd = synth_state_dict(0), synth_speaker_encoder(5)[0]
    tol = 1e-4 if precision == "fp32" else 1e-2
    k = 0
    for si, s in enumerate(sources):
        for ti, t in enumerate(targets):
            assert posts[k].shape == (80, 1 + len(s) // 275)
            if (si, ti) in ((0, 0), (1, 1), (2, 0)):                           # the oracle chain takes seconds per pair
                ref = _oracle_convert(sd, se_sd, s, s, t)
                err = float((posts[k].cpu() - ref).abs().max())
                print(f"{precision} pair ({si},{ti}): max-abs {err:.2e}")
                assert err <= tol
            k += 1
    # one conversion through `convert` is the same row
    one = vc.convert(sources[1], targets[1], vocode=False)
    assert float((one - posts[3]).abs().max()) <= (1e-6 if precision == "fp32" else 1e-2)


def test_embed_utterance_matches_oracle_and_batches():
    ae, se = _stack("fp32")
    se_sd = synth_speaker_encoder(5)[0]
    wavs = [_signal(40000, 4), _signal(52000, 5), _signal(20000, 6)]
    both = se.embed_utterances(wavs)
    for i, w in enumerate(wavs):
        ref = _oracle_embed(se_sd, w)
        alone, partials, slices = se.embed_utterance(w, return_partials=True)
        assert float((alone - ref).abs().max()) <= 1e-4 and float((both[i].cpu() - alone).abs().max()) <= 1e-5
        assert partials.shape[0] == len(slices) and slices[0] == slice(0, 25600)
    whole = se.embed_utterance(wavs[0], using_partials=False)
    assert whole.shape == (256,) and abs(float(whole.norm()) - 1) <= 1e-5


def test_cut_conversion_is_batch_forward():
    """`cut=True` (reference `:161-169`): chunks of the source as one batch, overlap-averaged."""
    from autovc_b200 import audio
    from autovc_b200.voice_converter import WavVoiceConverter
    ae, se = _stack("fp32")
    vc = WavVoiceConverter(ae, se, device="cuda")
    src, trg = _signal(60000, 7), _signal(40000, 8)
    post = vc.convert(src, trg, cut=True, vocode=False, partial_utterance_n_frames=64)
    chunks = torch.stack(audio.mel_spec_auto_encoder(src, cut=True, partial_utterance_n_frames=64))
    c_s, c_t = se.embed_utterance(src)[None].cuda(), se.embed_utterance(trg)[None].cuda()
    ref = ae.batch_forward(chunks, c_s, c_t, overlap=0.5)
    assert post.shape == ref.shape == (80, 64 + (chunks.shape[0] - 1) * 32)
    assert float((post - ref).abs().max()) <= 1e-6


def test_convert_to_waveform():
    """With a vocoder the result is what `WaveRNN.generate` makes of the converted mel (same random stream)."""
    import autovc_b200
    from autovc_b200.voice_converter import WavVoiceConverter
    ae, se = _stack("bf16")
    voc = autovc_b200.WaveRNN(device="cuda")
    voc.load_state_dict(synth_wavernn_state_dict(11))
    vc = WavVoiceConverter(ae, se, voc, device="cuda")
    src, trg = _signal(275 * 40, 9), _signal(40000, 4)
    post = vc.convert(src, trg, vocode=False)
    args = dict(batched=True, target=1500, overlap=100)
    steps = 0
    voc.generate(post.unsqueeze(0), **args); rows, steps = voc.last_rows, voc.last_steps
    U = synth_uniforms(steps, rows, 17)
    wav = vc.convert(src, trg, vocoder_args=dict(uniforms=U, **args))
    ref = voc.generate(post.unsqueeze(0), uniforms=U, **args)
    assert wav.dtype == np.float32 and wav.shape == ref.shape and np.isfinite(wav).all()
    assert np.array_equal(wav, ref)
    waves = vc.convert_multiple([src, src[: 275 * 33]], [trg], vocoder_args=dict(target=1500, overlap=100))
    assert len(waves) == 2 and all(np.isfinite(w).all() for w in waves)
