"""Parity tests proper (need a B200): the CUDA path, called through the custom op / C ABI, against
(1) the golden vectors generated from the unmodified reference, (2) the CPU oracle on the same seeded inputs,
(3) size-independent properties at BASELINE.json's full sizes.

Tolerances (north-star): fp32 validation mode max-abs <= 1e-4; bf16 mode max-abs(mel_outputs_postnet) <= 1e-2
and cosine(codes) >= 0.999 per utterance."""
import numpy as np
import pytest
import torch

from autovc_b200.synth import synth_inputs, synth_state_dict
from conftest import load_golden

pytestmark = pytest.mark.gpu

FP32_TOL = 1e-4
BF16_POST_TOL = 1e-2
BF16_COS = 0.999

DEFAULT_CASES = ["synth_b2_t128", "synth_b2_t100", "synth_b1_t33", "synth_b1_t32", "synth_b5_t64_w3",
                 "synth_b3_t400", "c1_samples_b2_t128"]


def _model(precision, wseed=0, gain=1.0):
    import autovc_b200
    m = autovc_b200.AutoEncoder(device="cuda", verbose=False, precision=precision)
    m.load_state_dict(synth_state_dict(wseed, recurrent_gain=gain))
    return m


_MODELS = {}


def model(precision, wseed=0, gain=1.0):
    key = (precision, wseed, gain)
    if key not in _MODELS:
        _MODELS[key] = _model(*key)
    return _MODELS[key]


def _case(name):
    g = load_golden(name)
    x, c_org, c_trg = synth_inputs(int(g["B"]), int(g["T"]), seed=int(g["xseed"]))
    if "x" in g.files:
        x = torch.from_numpy(g["x"])
    return g, x.cuda(), c_org.cuda(), c_trg.cuda()


@pytest.mark.parametrize("name", DEFAULT_CASES + ["stress_b2_t128_gain2p5"])
def test_fp32_mode_matches_reference_golden(name):
    g, x, c_org, c_trg = _case(name)
    m = model("fp32", int(g["wseed"]), float(g["gain"]))
    mel, post, codes = m(x, c_org, c_trg)
    torch.cuda.synchronize()
    assert mel.shape == x.shape and post.shape == x.shape and codes.shape == g["codes"].shape
    assert float((mel.cpu() - torch.from_numpy(g["mel"])).abs().max()) <= FP32_TOL
    assert float((post.cpu() - torch.from_numpy(g["mel_post"])).abs().max()) <= FP32_TOL
    assert float((codes.cpu() - torch.from_numpy(g["codes"])).abs().max()) <= FP32_TOL
    only = m(x, c_org, None)
    assert torch.equal(only, codes)


@pytest.mark.parametrize("name", DEFAULT_CASES)
def test_bf16_mode_matches_reference_golden(name):
    g, x, c_org, c_trg = _case(name)
    m = model("bf16", int(g["wseed"]), float(g["gain"]))
    mel, post, codes = m(x, c_org, c_trg)
    torch.cuda.synchronize()
    err_mel = float((mel.cpu() - torch.from_numpy(g["mel"])).abs().max())
    err_post = float((post.cpu() - torch.from_numpy(g["mel_post"])).abs().max())
    cos = float(torch.nn.functional.cosine_similarity(codes.cpu(), torch.from_numpy(g["codes"]), dim=1).min())
    print(f"{name}: bf16 mel {err_mel:.2e} post {err_post:.2e} cos {cos:.6f}")
    assert err_post <= BF16_POST_TOL and err_mel <= BF16_POST_TOL and cos >= BF16_COS
    only = m(x, c_org, None)
    assert torch.equal(only, codes)


def test_stress_weights_bf16_stays_close():
    """Recurrent gain 2.5 pushes the gates into saturation; bf16 is outside the 1e-2 budget by design here
    (precision) - just bound it so that a gross error is caught."""
    g, x, c_org, c_trg = _case("stress_b2_t128_gain2p5")
    m = model("bf16", int(g["wseed"]), float(g["gain"]))
    mel, post, codes = m(x, c_org, c_trg)
    assert float((post.cpu() - torch.from_numpy(g["mel_post"])).abs().max()) <= 0.25
    assert float(torch.nn.functional.cosine_similarity(codes.cpu(), torch.from_numpy(g["codes"]), dim=1).min()) >= 0.99


@pytest.mark.parametrize("precision", ["fp32", "bf16"])
def test_t_shorter_than_freq_raises_value_error(precision):
    m = model(precision)
    x, c_org, c_trg = synth_inputs(1, 31)
    with pytest.raises(ValueError):
        m(x.cuda(), c_org.cuda(), c_trg.cuda())


@pytest.mark.parametrize("precision,tol", [("fp32", FP32_TOL), ("bf16", BF16_POST_TOL)])
def test_batch_forward_overlap_average_golden(precision, tol):
    g = load_golden("batch_forward_m3_n128")
    m = model(precision, int(g["wseed"]))
    x, c_org, c_trg = synth_inputs(3, 128, seed=int(g["xseed"]))
    out = m.batch_forward(x.cuda(), c_org[:1].cuda(), c_trg[:1].cuda(), overlap=0.5)
    assert out.shape == (80, 256)
    assert float((out.cpu() - torch.from_numpy(g["out"])).abs().max()) <= tol


@pytest.mark.parametrize("precision,tol", [("fp32", FP32_TOL), ("bf16", BF16_POST_TOL)])
def test_against_live_oracle_long_utterance(precision, tol):
    """T=1024 (config 3's length; 4097 dependent recurrent steps), B=2, against the CPU oracle run here."""
    from oracle import autovc_oracle as O
    sd = synth_state_dict(0)
    x, c_org, c_trg = synth_inputs(2, 1024, seed=21)
    ref_mel, ref_post, ref_codes = O.generator_forward(sd, x, c_org, c_trg)
    mel, post, codes = model(precision)(x.cuda(), c_org.cuda(), c_trg.cuda())
    assert float((post.cpu() - ref_post).abs().max()) <= tol
    assert float((mel.cpu() - ref_mel).abs().max()) <= tol
    assert float(torch.nn.functional.cosine_similarity(codes.cpu(), ref_codes, dim=1).min()) >= BF16_COS


@pytest.mark.parametrize("precision", ["fp32", "bf16"])
def test_full_size_batch_rows_are_independent(precision):
    """BASELINE config 2 (B=256,T=128): every utterance is independent, so rows computed inside the
    big batch equal the same rows computed alone, and a sample of rows matches the oracle."""
    from oracle import autovc_oracle as O
    m = model(precision)
    x, c_org, c_trg = synth_inputs(256, 128, seed=3)
    mel, post, codes = m(x.cuda(), c_org.cuda(), c_trg.cuda())
    rows = [0, 77, 255]
    idx = torch.tensor(rows)
    mel_s, post_s, codes_s = m(x[idx].cuda(), c_org[idx].cuda(), c_trg[idx].cuda())
    assert float((post[idx.cuda()] - post_s).abs().max()) <= 1e-5
    assert float((codes[idx.cuda()] - codes_s).abs().max()) <= 1e-5
    ref_mel, ref_post, ref_codes = O.generator_forward(synth_state_dict(0), x[idx], c_org[idx], c_trg[idx])
    tol = FP32_TOL if precision == "fp32" else BF16_POST_TOL
    assert float((post_s.cpu() - ref_post).abs().max()) <= tol
    assert torch.isfinite(post).all() and torch.isfinite(mel).all()


def test_training_shape_forward_pair():
    """BASELINE config 5's two passes (reference model.py:306 then :201) at a reduced batch: pass 1 forward(X,c,c),
    pass 2 encoder-only on mel_outputs_postnet."""
    from oracle import autovc_oracle as O
    m = model("fp32")
    x, c_org, _ = synth_inputs(4, 128, seed=8, same_speaker=True)
    mel, post, codes = m(x.cuda(), c_org.cuda(), c_org.cuda())
    codes2 = m(post, c_org.cuda(), None)
    sd = synth_state_dict(0)
    r_mel, r_post, r_codes = O.generator_forward(sd, x, c_org, c_org)
    r_codes2 = O.generator_forward(sd, r_post, c_org, None)
    assert float((codes2.cpu() - r_codes2).abs().max()) <= FP32_TOL
    assert float((post.cpu() - r_post).abs().max()) <= FP32_TOL


def test_outputs_do_not_alias_cached_plans():
    """The plan cache holds workspace addresses only: fresh output tensors of later calls must be written even when
    earlier outputs are still alive (regression: channel-first output pointers were once cached in the GEMM plans)."""
    m = model("bf16")
    x, c_org, c_trg = (t.cuda() for t in synth_inputs(2, 128, seed=1234))
    keep = [m(x, c_org, c_trg) for _ in range(3)]       # all outputs alive -> three distinct allocations
    torch.cuda.synchronize()
    for mel, post, codes in keep[1:]:
        assert torch.equal(post, keep[0][1]) and torch.equal(mel, keep[0][0]) and torch.equal(codes, keep[0][2])
    assert float(keep[2][1].abs().max()) > 1.0


def test_forward_into_preallocated_outputs():
    """`out=` (extension used by HostPipeline so that its steady state never touches the allocator) fills caller tensors
    with exactly what the allocating call returns, and rejects tensors of the wrong shape."""
    m = model("bf16")
    x, c_org, c_trg = (t.cuda() for t in synth_inputs(3, 100, seed=4))
    mel, post, codes = m(x, c_org, c_trg)
    out = (torch.full_like(mel, float("nan")), torch.full_like(post, float("nan")), torch.full_like(codes, float("nan")))
    res = m(x, c_org, c_trg, out=out)
    assert res[0] is out[0] and torch.equal(out[0], mel) and torch.equal(out[1], post) and torch.equal(out[2], codes)
    with pytest.raises(RuntimeError):
        m(x, c_org, c_trg, out=(out[0][:2], out[1], out[2]))


@pytest.mark.parametrize("B,T", [(300, 64), (130, 96), (97, 33)])
def test_batches_beyond_one_recurrence_launch(B, T):
    """B > 256 runs the recurrences as several co-resident launches (256 + 44 rows), B = 130 takes 64-wide tiles for lstm2 and
    32-wide ones for lstm1, B = 97 is ragged in both: every row equals the same row computed in a small batch, and a sample
    of rows matches the oracle."""
    from oracle import autovc_oracle as O
    m = model("bf16")
    x, c_org, c_trg = synth_inputs(B, T, seed=B)
    mel, post, codes = m(x.cuda(), c_org.cuda(), c_trg.cuda())
    rows = [0, B // 2, 255 if B > 256 else B - 2, B - 1]
    idx = torch.tensor(rows)
    mel_s, post_s, codes_s = m(x[idx].cuda(), c_org[idx].cuda(), c_trg[idx].cuda())
    assert float((post[idx.cuda()] - post_s).abs().max()) <= 1e-5
    assert float((codes[idx.cuda()] - codes_s).abs().max()) <= 1e-5
    ref_mel, ref_post, ref_codes = O.generator_forward(synth_state_dict(0), x[idx], c_org[idx], c_trg[idx])
    assert float((post_s.cpu() - ref_post).abs().max()) <= BF16_POST_TOL
    assert torch.isfinite(post).all()


# ---- full-size parity (BASELINE.json configs 2, 3, 5) against the oracle on sampled rows ----------------------------
def _check_rows(precision, got, ref, rows_note):
    mel, post, codes = got
    r_mel, r_post, r_codes = ref
    tol = FP32_TOL if precision == "fp32" else BF16_POST_TOL
    e_mel, e_post = float((mel.cpu() - r_mel).abs().max()), float((post.cpu() - r_post).abs().max())
    cos = float(torch.nn.functional.cosine_similarity(codes.cpu(), r_codes, dim=1).min())
    print(f"{rows_note}[{precision}]: mel {e_mel:.2e} post {e_post:.2e} cos {cos:.6f}")
    assert e_mel <= tol and e_post <= tol and cos >= BF16_COS
    if precision == "fp32":
        assert float((codes.cpu() - r_codes).abs().max()) <= FP32_TOL


@pytest.mark.parametrize("precision", ["fp32", "bf16"])
def test_config2_sixteen_rows_match_oracle(precision):
    """Config 2 (B=256, T=128): 16 seeded-random rows of the full batch against the oracle (rows are independent, so the
    oracle runs on those rows alone)."""
    from oracle import autovc_oracle as O
    m = model(precision)
    x, c_org, c_trg = synth_inputs(256, 128, seed=1234)
    mel, post, codes = m(x.cuda(), c_org.cuda(), c_trg.cuda())
    idx = torch.from_numpy(np.sort(np.random.default_rng(7).choice(256, size=16, replace=False)))
    ref = O.generator_forward(synth_state_dict(0), x[idx], c_org[idx], c_trg[idx])
    _check_rows(precision, (mel[idx.cuda()], post[idx.cuda()], codes[idx.cuda()]), ref, "config2 rows " + str(idx.tolist()))


@pytest.mark.parametrize("precision", ["fp32", "bf16"])
def test_config3_full_batch_rows_match_oracle(precision):
    """Config 3 at its full size (B=64, T=1024: the 32-utterance lstm2 tiles, 2048 dependent lstm2 steps): rows at both
    ends of both tiles against the oracle."""
    from oracle import autovc_oracle as O
    m = model(precision)
    x, c_org, c_trg = synth_inputs(64, 1024, seed=33)
    mel, post, codes = m(x.cuda(), c_org.cuda(), c_trg.cuda())
    idx = torch.tensor([0, 31, 32, 63])
    ref = O.generator_forward(synth_state_dict(0), x[idx], c_org[idx], c_trg[idx])
    _check_rows(precision, (mel[idx.cuda()], post[idx.cuda()], codes[idx.cuda()]), ref, "config3 rows 0/31/32/63")


def test_config5_bf16_pair_full_batch_rows_match_oracle():
    """Config 5's per-GPU share (B=256, T=128, eval-mode BatchNorm), bf16: pass 1 forward(X,c,c) and pass 2
    forward(mel_post,c,None) on four sampled rows against the oracle's pair on those rows."""
    from oracle import autovc_oracle as O
    m = model("bf16")
    x, c_org, _ = synth_inputs(256, 128, seed=58)
    mel, post, codes = m(x.cuda(), c_org.cuda(), c_org.cuda())
    codes2 = m(post, c_org.cuda(), None)
    idx = torch.tensor([0, 100, 200, 255])
    sd = synth_state_dict(0)
    r_mel, r_post, r_codes = O.generator_forward(sd, x[idx], c_org[idx], c_org[idx])
    r_codes2 = O.generator_forward(sd, r_post, c_org[idx], None)
    _check_rows("bf16", (mel[idx.cuda()], post[idx.cuda()], codes[idx.cuda()]), (r_mel, r_post, r_codes), "config5 pass 1")
    cos2 = float(torch.nn.functional.cosine_similarity(codes2[idx.cuda()].cpu(), r_codes2, dim=1).min())
    print(f"config5 pass 2: cos(codes) {cos2:.6f}")
    assert cos2 >= BF16_COS


def test_recurrence_exchange_is_bitwise_reproducible_under_stress():
    """The in-band h exchange (ll_lstm.cu: the data is the flag) claims a timing-independent accumulation order: 200
    forwards of config 3 (B=64, T=1024: 0.8 M exchanged step tiles in all) must give bitwise-identical outputs."""
    m = model("bf16")
    x, c_org, c_trg = (t.cuda() for t in synth_inputs(64, 1024, seed=91))
    first = [t.clone() for t in m(x, c_org, c_trg)]
    out = tuple(torch.empty_like(t) for t in first)
    bad = 0
    for _ in range(200):
        m(x, c_org, c_trg, out=out)
        bad += int(not all(torch.equal(a, b) for a, b in zip(out, first)))
    assert bad == 0, f"{bad} of 200 runs differed"
    assert torch.isfinite(first[1]).all()


def test_plan_cache_keeps_alternating_shapes():
    """Alternating shapes and modes (length buckets, B=1 after a big batch, the encode-only pass) must not rebuild tensor
    maps: each (B, T, mode) is built once and results stay identical."""
    import autovc_b200
    m = autovc_b200.AutoEncoder(device="cuda", verbose=False, precision="bf16")
    m.load_state_dict(synth_state_dict(0))
    big = tuple(t.cuda() for t in synth_inputs(64, 128, seed=5))      # sizes the workspace first: its address then stays put
    small = tuple(t.cuda() for t in synth_inputs(1, 128, seed=6))
    rag = tuple(t.cuda() for t in synth_inputs(3, 100, seed=7))
    want = [m(*big), m(*small), m(*rag), m(big[0], big[1], None)]
    eng = m._engine(torch.device("cuda", 0))
    builds = eng.plan_builds
    for _ in range(3):
        got = [m(*big), m(*small), m(*rag), m(big[0], big[1], None)]
        for g, w in zip(got, want):
            for a, b in zip(g if isinstance(g, tuple) else (g,), w if isinstance(w, tuple) else (w,)):
                assert torch.equal(a, b)
    assert eng.plan_builds == builds, (eng.plan_builds, builds)


def test_weight_changes_reach_the_packed_copy():
    """In-place updates through autograd-visible ops are picked up automatically; writes through `.data` (the reference's
    own load_flattened_params, model.py:385-392) need invalidate() - and this class's load_flattened_params does it."""
    import autovc_b200
    m = autovc_b200.AutoEncoder(device="cuda", verbose=False, precision="fp32")
    m.load_state_dict(synth_state_dict(0))
    x, c_org, c_trg = (t.cuda() for t in synth_inputs(2, 64, seed=2))
    base = m(x, c_org, c_trg)[1].clone()
    w = m.decoder.linear_projection.linear_layer.bias
    saved = w.detach().clone()
    with torch.no_grad():
        w.add_(0.5)
    shifted = m(x, c_org, c_trg)[1]
    assert float((shifted - base).abs().max()) > 0.1
    w.data.copy_(saved)                     # invisible to the version counter ...
    assert torch.equal(m(x, c_org, c_trg)[1], shifted)
    m.invalidate()                          # ... until the caller says so
    assert float((m(x, c_org, c_trg)[1] - base).abs().max()) <= 1e-6
    flat = m.flatten_params()
    m.load_flattened_params(flat * 0 + flat)
    assert float((m(x, c_org, c_trg)[1] - base).abs().max()) <= 1e-6


def test_caller_device_is_restored():
    """avc_forward switches to the context's device for its launches and must leave the caller's current device alone."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    import autovc_b200
    m = autovc_b200.AutoEncoder(device="cuda:1", verbose=False, precision="bf16")
    m.load_state_dict(synth_state_dict(0))
    x, c_org, c_trg = (t.to("cuda:1") for t in synth_inputs(2, 64, seed=2))
    torch.cuda.set_device(0)
    m(x, c_org, c_trg)
    probe = torch.empty(1, device="cuda")        # allocated on the CUDA runtime's current device
    assert torch.cuda.current_device() == 0 and probe.device.index == 0


@pytest.mark.parametrize("B,T", [(3, 33), (2, 128), (5, 100), (1, 32)])
def test_fused_speaker_concat_equals_materialised_concat(B, T, monkeypatch):
    """Encoder conv0 with the speaker concat fused away (mel-only convolution + per-utterance edge-class bias vectors,
    reference encoder.py:65-70) against the same layer over the materialised [B,T,336] operand (AVC_NO_FUSED_CONCAT=1):
    the two differ only in where the fp32 partial sums are added, and both stay within the oracle tolerance - including
    the first / last two frames of every utterance, whose windows reach into the zero padding."""
    from oracle import autovc_oracle as O
    m = model("bf16")
    x, c_org, c_trg = synth_inputs(B, T, seed=700 + T)
    fused = [t.clone() for t in m(x.cuda(), c_org.cuda(), c_trg.cuda())]
    monkeypatch.setenv("AVC_NO_FUSED_CONCAT", "1")
    plain = [t.clone() for t in m(x.cuda(), c_org.cuda(), c_trg.cuda())]
    monkeypatch.delenv("AVC_NO_FUSED_CONCAT")
    ref = O.generator_forward(synth_state_dict(0), x, c_org, c_trg)
    d_codes = float((fused[2] - plain[2]).abs().max())
    d_post = float((fused[1] - plain[1]).abs().max())
    print(f"fused vs materialised B={B} T={T}: codes {d_codes:.2e} post {d_post:.2e}")
    assert d_codes <= 2e-2 and d_post <= BF16_POST_TOL
    for got in (fused, plain):
        _check_rows("bf16", got, ref, f"concat B={B} T={T}")
    # edge frames of the codes' source: the encoder output at t = 0 feeds the first backward code
    n_f = T // 32
    assert float((fused[2][:, 32 * n_f:32 * n_f + 32].cpu() - ref[2][:, 32 * n_f:32 * n_f + 32]).abs().max()) <= 2e-2


@pytest.mark.parametrize("B,T,mode", [(3, 64, "forward"), (5, 100, "forward"), (2, 33, "encode"), (70, 40, "forward")])
def test_workspace_and_outputs_stay_inside_their_bounds(B, T, mode):
    """compute-sanitizer is closed on this GPU pool (profiles/r2_sanitizer.md), so out-of-bounds WRITES are hunted with guard
    bands: the workspace and every output of avc_forward / avc_encode sit between 64 KiB bands of a known pattern inside
    one allocation, called through the C ABI with exactly avc_workspace_bytes; afterwards every band must be untouched and
    the results must equal those of the normal path."""
    import ctypes
    from autovc_b200 import _lib as L
    m = model("bf16")
    eng = m._engine(torch.device("cuda", 0))
    x, c_org, c_trg = (t.cuda().contiguous() for t in synth_inputs(B, T, seed=900 + B))
    want = m(x, c_org, c_trg) if mode == "forward" else (None, None, m(x, c_org, None))
    G = 64 << 10
    need = eng.workspace_bytes(B, T)
    n_codes = 32 * (T // 32 + (T + 31) // 32)
    sizes = [need, B * 80 * T * 4, B * 80 * T * 4, B * n_codes * 4]          # workspace, mel, mel_post, codes
    offs, total = [], G
    for sz in sizes:
        offs.append(total)
        total += (sz + 255) // 256 * 256 + G
    arena = torch.full((total,), 0xA5, dtype=torch.uint8, device="cuda")
    base = arena.data_ptr()
    assert base % 256 == 0
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    if mode == "forward":
        rc = L.lib().avc_forward(eng._h, x.data_ptr(), c_org.data_ptr(), c_trg.data_ptr(), B, T, base + offs[1], base + offs[2],
                                 base + offs[3], base + offs[0], need, stream)
    else:
        rc = L.lib().avc_encode(eng._h, x.data_ptr(), c_org.data_ptr(), B, T, base + offs[3], base + offs[0], need, stream)
    assert rc == L.AVC_OK, L.last_error(eng._h)
    torch.cuda.synchronize()
    bands = [(0, G)] + [(o + (sz + 255) // 256 * 256, o + (sz + 255) // 256 * 256 + G) for o, sz in zip(offs, sizes)]
    for lo, hi in bands:
        assert bool((arena[lo:hi] == 0xA5).all()), f"guard band [{lo}, {hi}) was written"
    # the slack between an output's end and its 256-byte rounding must be untouched too
    for o, sz in zip(offs[1:], sizes[1:]):
        end = o + (sz + 255) // 256 * 256
        assert bool((arena[o + sz:end] == 0xA5).all())
    codes = arena[offs[3]:offs[3] + sizes[3]].view(torch.float32).view(B, n_codes)
    assert torch.equal(codes, want[2])
    if mode == "forward":
        post = arena[offs[2]:offs[2] + sizes[2]].view(torch.float32).view(B, 80, T)
        mel = arena[offs[1]:offs[1] + sizes[1]].view(torch.float32).view(B, 80, T)
        assert torch.equal(post, want[1]) and torch.equal(mel, want[0])
    else:                                           # encode-only must not touch the mel outputs' slots at all
        assert bool((arena[offs[1]:offs[1] + sizes[1]] == 0xA5).all()) and bool((arena[offs[2]:offs[2] + sizes[2]] == 0xA5).all())


def test_repeated_calls_replay_one_cuda_graph(monkeypatch):
    """A call that repeats an earlier call's pointers runs as ONE graph launch (built on the second such call) with results
    bitwise equal to the plain launch sequence (AVC_NO_GRAPH=1); a different pointer set gets its own graph; profiling mode
    and a reloaded state dict fall back to / rebuild the plain sequence."""
    import autovc_b200
    sd = synth_state_dict(0)
    x, c_org, c_trg = (t.cuda() for t in synth_inputs(5, 100, seed=31))
    monkeypatch.setenv("AVC_NO_GRAPH", "1")
    plain_m = autovc_b200.AutoEncoder(device="cuda", verbose=False, precision="bf16")
    plain_m.load_state_dict(sd)
    want = [t.clone() for t in plain_m(x, c_org, c_trg)]
    want_codes = plain_m(x, c_org, None).clone()
    assert plain_m._engine(x.device).graph_replays == 0
    monkeypatch.delenv("AVC_NO_GRAPH")
    m = autovc_b200.AutoEncoder(device="cuda", verbose=False, precision="bf16")
    m.load_state_dict(sd)
    eng = m._engine(x.device)
    out_a = tuple(torch.empty_like(t) for t in want)
    out_b = tuple(torch.empty_like(t) for t in want)
    for i in range(5):
        for out in (out_a, out_b):
            for t in out:
                t.fill_(float("nan"))
            m(x, c_org, c_trg, out=out)
            assert all(torch.equal(a, b) for a, b in zip(out, want)), f"call {i}"
    # out_a: call 0 builds the plans, call 1 is the first sighting of its pointers, calls 2-4 run as a graph (3);
    # out_b: call 0 is the first sighting, calls 1-4 run as a graph (4)
    assert eng.graph_replays == 7
    n = eng.graph_replays
    for _ in range(4):                                    # fresh output tensors: torch's allocator hands the same blocks back
        assert torch.equal(m(x, c_org, None), want_codes)
    assert eng.graph_replays > n
    n, launches = eng.graph_replays, eng.last_launch_count
    assert launches >= 6
    eng.set_profiling(True)
    m(x, c_org, c_trg, out=out_a)
    assert eng.graph_replays == n and sum(eng.profile().values()) > 0
    eng.set_profiling(False)
    m.load_state_dict(sd)                                 # weights re-packed: plans and graphs are rebuilt, results unchanged
    for _ in range(3):
        m(x, c_org, c_trg, out=out_a)
    assert all(torch.equal(a, b) for a, b in zip(out_a, want))


def test_forward_inside_a_callers_own_graph_capture():
    """When the caller is capturing (torch.cuda.graph) the library must not start a capture of its own: the plain sequence is
    recorded into the caller's graph, and replaying that graph reproduces the eager result."""
    m = model("bf16")
    x, c_org, c_trg = (t.cuda() for t in synth_inputs(3, 64, seed=41))
    want = [t.clone() for t in m(x, c_org, c_trg)]
    out = tuple(torch.empty_like(t) for t in want)
    for _ in range(3):
        m(x, c_org, c_trg, out=out)                        # the library's own graph for this pointer set exists by now
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    side = torch.cuda.Stream()
    with torch.cuda.stream(side):
        with torch.cuda.graph(g, stream=side):
            m(x, c_org, c_trg, out=out)
    for t in out:
        t.zero_()
    g.replay()
    torch.cuda.synchronize()
    assert all(torch.equal(a, b) for a, b in zip(out, want))


def test_nan_input_reaches_the_outputs_without_stalling_the_exchange():
    """The recurrence's exchange ring uses bit 14 of every bf16 h as its epoch tag (|h| <= 1 keeps it free); a NaN h would carry
    that bit itself.  A NaN frame in one utterance must neither stall / trap the poll loops nor touch the other utterances, and
    must still show up as NaN in that utterance's outputs (the untagged output copy keeps it)."""
    m = model("bf16")
    x, c_org, c_trg = (t.cuda() for t in synth_inputs(3, 96, seed=51))
    clean = [t.clone() for t in m(x, c_org, c_trg)]
    xn = x.clone()
    xn[1, :, 40] = float("nan")
    mel, post, codes = m(xn, c_org, c_trg)
    torch.cuda.synchronize()
    assert torch.isnan(post[1]).any() and torch.isnan(mel[1]).any()
    for r in (0, 2):
        assert torch.equal(post[r], clean[1][r]) and torch.equal(codes[r], clean[2][r])
    again = m(x, c_org, c_trg)                         # and the context is still healthy
    assert all(torch.equal(a, b) for a, b in zip(again, clean))


@pytest.mark.parametrize("B,T", [(7, 57), (33, 45), (2, 257), (17, 96), (1, 129), (40, 32), (129, 64), (65, 33)])
def test_ragged_shapes_match_oracle(B, T):
    """Ragged (B, T) pairs - partial 32-utterance tiles, one and two sub-tiles per CTA pair (B = 129 crosses the switch), T not a
    multiple of 32 or of the GEMM tile, T = freq - in bf16 against the oracle on up to four sampled rows, and in fp32 on two."""
    from oracle import autovc_oracle as O
    sd = synth_state_dict(0)
    x, c_org, c_trg = synth_inputs(B, T, seed=1000 + 7 * B + T)
    rows = sorted({0, B // 3, (2 * B) // 3, B - 1})
    idx = torch.tensor(rows)
    ref = O.generator_forward(sd, x[idx], c_org[idx], c_trg[idx])
    mel, post, codes = model("bf16")(x.cuda(), c_org.cuda(), c_trg.cuda())
    assert codes.shape == (B, 32 * (T // 32 + (T + 31) // 32))
    _check_rows("bf16", (mel[idx.cuda()], post[idx.cuda()], codes[idx.cuda()]), ref, f"ragged B={B} T={T}")
    assert torch.isfinite(post).all()
    two = idx[:2]
    mel32, post32, codes32 = model("fp32")(x[two].cuda(), c_org[two].cuda(), c_trg[two].cuda())
    _check_rows("fp32", (mel32, post32, codes32), tuple(r[:2] for r in ref), f"ragged fp32 B={B} T={T}")


@pytest.mark.parametrize("B,T", [(9, 257), (257, 128), (300, 64), (129, 45), (31, 128), (160, 96), (17, 200), (96, 100), (8, 129), (65, 33)])
def test_rows_are_bitwise_independent_of_the_batch_across_kernel_variants(B, T):
    """A row's bf16 result is bit for bit the same alone (one sub-tile, 16 group warps and N = 16 MMAs in the H = 1024 recurrence), in
    a batch of nine, and inside the full batch (two sub-tiles from B = 129, ragged last tiles, several launches from B = 257): the
    accumulation order of the recurrence and of the GEMMs does not depend on the kernel variant the batch size selects.
    (`tools/fuzz_rows.py` runs the same check over random shapes.)"""
    m = model("bf16")
    x, c_org, c_trg = (t.cuda() for t in synth_inputs(B, T, seed=7000 + 3 * B + T))
    full = [t.clone() for t in m(x, c_org, c_trg)]
    for r in sorted({0, B // 2, B - 1}):
        one = m(x[r:r + 1], c_org[r:r + 1], c_trg[r:r + 1])
        assert all(torch.equal(a[0], f[r]) for a, f in zip(one, full)), f"row {r} alone differs from row {r} of B={B}"
    k = min(B, 9)
    sub = m(x[:k], c_org[:k], c_trg[:k])
    assert all(torch.equal(a, f[:k]) for a, f in zip(sub, full))
    assert torch.isfinite(full[1]).all()
