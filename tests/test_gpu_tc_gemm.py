"""Per-layer check of the tcgen05/TMA implicit-GEMM kernel against a plain fp32 PyTorch restatement of the same
op (same bf16/fp16-rounded operands), through the C ABI's avc_debug_gemm.  Needs a B200."""
import pytest
import torch

from autovc_b200.synth import synth_state_dict

pytestmark = pytest.mark.gpu

# layer, Cin (row length of A), operand dtype, activation (0 none, 1 relu, 2 tanh), channel-first output?
LAYERS = [
    ("enc.conv0", 336, torch.bfloat16, 1, False),
    ("enc.conv1", 512, torch.bfloat16, 1, False),
    ("enc.lstm.l0.ih", 512, torch.bfloat16, 0, False),
    ("enc.lstm.l1.ih", 64, torch.bfloat16, 0, False),
    ("dec.lstm1.ih", 320, torch.bfloat16, 0, False),
    ("dec.lstm2.l0.ih", 512, torch.bfloat16, 0, False),
    ("dec.lstm2.l1.ih", 1024, torch.bfloat16, 0, False),
    ("dec.linear", 1024, torch.bfloat16, 0, True),
    ("post.conv0", 80, torch.float16, 2, False),
    ("post.conv2", 512, torch.float16, 2, False),
    ("post.conv4", 512, torch.float16, 0, True),
]
SHAPES = [(2, 128), (3, 100), (5, 33), (2, 400), (40, 128)]


@pytest.fixture(scope="module")
def engine():
    from autovc_b200.engine import Engine
    eng = Engine((32, 256, 512, 32), 0, "bf16")
    eng.load_state_dict(synth_state_dict(0))
    yield eng
    eng.close()


def reference(eng, layer, A, B, T, act, dtype):
    """fp32 restatement: unfold the taps explicitly, matmul with the packed weight (rounded like the kernel's)."""
    w = eng.packed(layer + ".w").cuda()
    b = eng.packed(layer + ".b").cuda()
    N = b.numel()
    Kp = w.numel() // N
    w = w.view(N, Kp).to(dtype).float()
    C = A.shape[1]
    a = A.float().view(B, T, C)
    if "conv" in layer:
        Cp = Kp // 5
        ap = torch.zeros(B, T + 4, Cp, device=A.device)
        ap[:, 2:T + 2, :C] = a
        cols = torch.cat([ap[:, k:k + T, :] for k in range(5)], dim=2)      # (B,T,5*Cp), k = tap*Cp + ci
    else:
        cols = torch.zeros(B, T, Kp, device=A.device)
        cols[:, :, :C] = a
    y = torch.matmul(cols.double(), w.double().t()).float() + b
    if act == 1:
        y = torch.relu(y)
    elif act == 2:
        y = torch.tanh(y)
    return y.view(B * T, N)


@pytest.mark.parametrize("layer,cin,dtype,act,chan_first", LAYERS)
@pytest.mark.parametrize("B,T", SHAPES)
def test_tc_gemm_layer(engine, layer, cin, dtype, act, chan_first, B, T):
    g = torch.Generator(device="cuda").manual_seed(B * 1000 + T)
    A = (torch.rand(B * T, cin, device="cuda", generator=g) * 2 - 1).to(dtype).contiguous()
    ref = reference(engine, layer, A, B, T, act, dtype)
    N = ref.shape[1]
    out_dtype = torch.float32 if "ih" in layer else dtype
    out = torch.full((B * T, N), float("nan"), device="cuda", dtype=out_dtype)
    out_t = res = None
    if chan_first:
        out_t = torch.full((B, N, T), float("nan"), device="cuda")
        res = torch.rand(B, N, T, device="cuda", generator=g) if layer == "post.conv4" else None
    engine.debug_gemm(layer, A, T, act, None if layer == "post.conv4" else out, True, out_t, res)
    torch.cuda.synchronize()
    tol = 2e-2 if out_dtype != torch.float32 else 2e-3     # output rounding of bf16 (2^-9 * |y|) dominates
    if layer != "post.conv4":
        err = float((out.float() - ref).abs().max())
        assert err <= tol * max(1.0, float(ref.abs().max())), f"{layer} B={B} T={T}: max err {err}"
    if chan_first:
        want = ref.view(B, T, N).transpose(1, 2) + (res if res is not None else 0)
        err = float((out_t - want).abs().max())
        assert err <= 2e-3 * max(1.0, float(ref.abs().max())), f"{layer} B={B} T={T}: channel-first max err {err}"


@pytest.mark.parametrize("layer,cin,dtype,act,chan_first", LAYERS[:3])
def test_simt_gemm_agrees_with_tc(engine, layer, cin, dtype, act, chan_first):
    B, T = 3, 100
    A = (torch.rand(B * T, cin, device="cuda") * 2 - 1).to(dtype).contiguous()
    ref = reference(engine, layer, A, B, T, act, dtype)
    out_dtype = torch.float32 if "ih" in layer else dtype
    out = torch.empty(B * T, ref.shape[1], device="cuda", dtype=out_dtype)
    engine.debug_gemm(layer, A, T, act, out, False)
    torch.cuda.synchronize()
    assert float((out.float() - ref).abs().max()) <= 2e-2 * max(1.0, float(ref.abs().max()))


@pytest.mark.parametrize("layer,cin,dtype,act", [("enc.conv1", 512, torch.bfloat16, 1), ("post.conv2", 512, torch.float16, 2)])
def test_tail_round_split_into_half_tiles(engine, layer, cin, dtype, act):
    """N=512 layers at B*T = 100 x 128 rows: 200 tiles of 128x256 on 148 CTAs leave a last round of 52 tiles, which the
    static schedule splits into 104 half-width tiles (tc_gemm.cu::decode_tile).  Results must not depend on the split."""
    import os
    B, T = 100, 128
    A = (torch.rand(B * T, cin, device="cuda") * 2 - 1).to(dtype).contiguous()
    ref = reference(engine, layer, A, B, T, act, dtype)
    out = torch.full((B * T, 512), float("nan"), device="cuda", dtype=dtype)
    engine.debug_gemm(layer, A, T, act, out, True)
    torch.cuda.synchronize()
    assert float((out.float() - ref).abs().max()) <= 2e-2 * max(1.0, float(ref.abs().max()))
    os.environ["AVC_GEMM_NO_TAIL_SPLIT"] = "1"
    try:
        out2 = torch.full_like(out, float("nan"))
        engine.debug_gemm(layer, A, T, act, out2, True)
        torch.cuda.synchronize()
    finally:
        del os.environ["AVC_GEMM_NO_TAIL_SPLIT"]
    assert torch.equal(out, out2)
