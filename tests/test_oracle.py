"""The oracle (oracle/autovc_oracle.py) against the known-answer vectors generated from the UNMODIFIED
reference module by oracle/make_golden.py (tests/golden/*.npz).  CPU only."""
import numpy as np
import pytest
import torch

from autovc_b200.synth import synth_inputs, synth_state_dict
from oracle import autovc_oracle as O
from conftest import load_golden

CASES = ["synth_b2_t128", "synth_b2_t100", "synth_b1_t33", "synth_b1_t32", "synth_b5_t64_w3",
         "stress_b2_t128_gain2p5", "c1_samples_b2_t128"]
FP32_TOL = 2e-5   # oracle (explicit matmuls) vs reference (oneDNN conv / ATen LSTM), both fp32


def _inputs(g):
    sd = synth_state_dict(int(g["wseed"]), recurrent_gain=float(g["gain"]))
    x, c_org, c_trg = synth_inputs(int(g["B"]), int(g["T"]), seed=int(g["xseed"]))
    if "x" in g.files:
        x = torch.from_numpy(g["x"])
    return sd, x, c_org, c_trg


@pytest.mark.parametrize("name", CASES)
def test_oracle_matches_reference_golden(name):
    g = load_golden(name)
    sd, x, c_org, c_trg = _inputs(g)
    mel, post, codes = O.generator_forward(sd, x, c_org, c_trg)
    assert mel.shape == (int(g["B"]), 80, int(g["T"]))
    assert float((mel - torch.from_numpy(g["mel"])).abs().max()) <= FP32_TOL
    assert float((post - torch.from_numpy(g["mel_post"])).abs().max()) <= FP32_TOL
    assert float((codes - torch.from_numpy(g["codes"])).abs().max()) <= FP32_TOL
    # c_trg=None returns the same codes (reference model.py:98-101)
    only = O.generator_forward(sd, x, c_org, None)
    assert torch.equal(only, codes)


def test_oracle_long_utterance_golden():
    g = load_golden("synth_b3_t400")
    sd, x, c_org, c_trg = _inputs(g)
    mel, post, codes = O.generator_forward(sd, x, c_org, c_trg)
    assert float((post - torch.from_numpy(g["mel_post"])).abs().max()) <= FP32_TOL
    assert codes.shape == (3, 32 * (12 + 13))


def test_oracle_intermediate_taps():
    g = load_golden("synth_b2_t128")
    sd, x, c_org, c_trg = _inputs(g)
    taps = {}
    O.generator_forward(sd, x, c_org, c_trg, taps=taps)
    # reference hooks captured BatchNorm outputs (pre-activation); oracle taps are post-activation
    ref = torch.from_numpy(g["tap.encoder.convbn2"])
    assert float((torch.clamp_min(ref, 0) - taps["encoder.conv2"][0, ::8, :]).abs().max()) <= FP32_TOL
    ref = torch.from_numpy(g["tap.postnet.convbn3"])
    assert float((torch.tanh(ref) - taps["postnet.conv3"][0, ::8, :]).abs().max()) <= FP32_TOL
    for key, tap in (("encoder.lstm", "encoder.lstm.l1"), ("decoder.lstm1", "decoder.lstm1.l0"), ("decoder.lstm2", "decoder.lstm2.l1")):
        ref = torch.from_numpy(g["tap." + key])
        assert float((ref - taps[tap][0, :, ::8]).abs().max()) <= FP32_TOL


def test_code_layout_forward_then_backward():
    """content_codes = all forward codes (t=31,63,..) then all backward codes (t=0,32,..)."""
    sd = synth_state_dict(0)
    x, c_org, _ = synth_inputs(1, 100)
    taps = {}
    codes = O.generator_forward(sd, x, c_org, None, taps=taps)
    y = taps["encoder.lstm.l1"]          # (1,100,64)
    assert codes.shape == (1, 32 * (3 + 4))
    for j, t in enumerate((31, 63, 95)):
        assert torch.equal(codes[0, 32 * j:32 * (j + 1)], y[0, t, :32])
    for j, t in enumerate((0, 32, 64, 96)):
        assert torch.equal(codes[0, 96 + 32 * j:96 + 32 * (j + 1)], y[0, t, 32:])


def test_t_shorter_than_freq_raises_like_reference():
    g = load_golden("batch_forward_m3_n128")
    assert str(g["t_lt_freq_raises"]) == "ValueError"
    sd = synth_state_dict(0)
    x, c_org, c_trg = synth_inputs(1, 31)
    with pytest.raises(ValueError):
        O.generator_forward(sd, x, c_org, c_trg)


def test_batch_forward_overlap_average_golden():
    g = load_golden("batch_forward_m3_n128")
    sd = synth_state_dict(int(g["wseed"]))
    x, c_org, c_trg = synth_inputs(3, 128, seed=int(g["xseed"]))
    out = O.batch_forward(sd, x, c_org[:1], c_trg[:1], overlap=0.5)
    assert out.shape == (80, 256)
    assert float((out - torch.from_numpy(g["out"])).abs().max()) <= FP32_TOL


def test_speaker_encoder_golden():
    g = load_golden("speaker_encoder_p8")
    from autovc_b200.synth import synth_speaker_encoder
    sd, parts = synth_speaker_encoder(int(g["seed"]))
    emb = O.speaker_encoder_forward(sd, parts)
    assert float((emb - torch.from_numpy(g["partial_embeds"])).abs().max()) <= FP32_TOL
    utt = O.embed_utterance_from_partials(sd, parts)
    assert float((utt - torch.from_numpy(g["embed"])).abs().max()) <= FP32_TOL


def test_bf16_error_budget_prediction():
    """CPU emulation of the product's operand rounding: bf16 everywhere except fp16 in the postnet keeps
    mel_outputs_postnet within the north-star 1e-2; bf16 in the postnet would not (precision)."""
    g = load_golden("synth_b2_t128")
    sd, x, c_org, c_trg = _inputs(g)
    f16 = lambda t: t.to(torch.float16).to(t.dtype)
    mixed = lambda t, tag: f16(t) if tag.startswith("postnet") else O.bf16_round(t)
    _, post, codes = O.generator_forward(sd, x, c_org, c_trg, quant=mixed)
    ref_post, ref_codes = torch.from_numpy(g["mel_post"]), torch.from_numpy(g["codes"])
    assert float((post - ref_post).abs().max()) < 6e-3
    assert float(torch.nn.functional.cosine_similarity(codes, ref_codes, dim=1).min()) > 0.9999
    _, post_bf16, _ = O.generator_forward(sd, x, c_org, c_trg, quant=O.bf16_round)
    assert float((post_bf16 - ref_post).abs().max()) > 1e-2
