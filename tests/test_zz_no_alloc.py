"""avc_se_forward_static (include/autovc_b200.h): the allocation-free variant of avc_se_forward, whose scratch is a static
device array inside the library.  It has no counterpart in the reference.

Called through the C ABI with caller-owned input / output tensors only, twice in a row on different inputs (the array keeps
the previous call's contents), and checked against the existing entry point avc_se_forward within the speaker-encoder
tests' bf16 tolerance (max-abs <= 1e-3 on the unit-norm embeddings; the two run the same kernels, so they agree far
closer).  The file name makes pytest collect it last."""
import ctypes

import pytest
import torch

from autovc_b200.synth import synth_speaker_encoder

pytestmark = pytest.mark.gpu

BF16_TOL = 1e-3        # tests/test_gpu_speaker_encoder.py


@pytest.fixture(scope="module")
def se_engine():
    from autovc_b200.engine import Engine
    sd, _ = synth_speaker_encoder(5)
    eng = Engine((32, 256, 512, 32), 0, "bf16")
    eng.load_state_dict(sd, speaker_encoder=True)
    yield eng
    eng.close()


def _static_forward(eng, partials, out):
    from autovc_b200 import _lib as L
    N, T, _ = partials.shape
    stream = ctypes.c_void_p(torch.cuda.current_stream(partials.device).cuda_stream)
    return L.lib().avc_se_forward_static(eng._h, partials.data_ptr(), N, T, out.data_ptr(), stream)


def test_static_scratch_entry_point_matches_workspace_entry_point(se_engine):
    from autovc_b200 import _lib as L
    g = torch.Generator().manual_seed(77)
    batches = [torch.randn(8, 160, 40, generator=g).cuda(), (1.5 * torch.randn(8, 160, 40, generator=g)).cuda(),
               torch.randn(3, 100, 40, generator=g).cuda()]
    for i, partials in enumerate(batches):                          # calls in a row on one shared array
        want = se_engine.se_forward(partials)                       # existing entry point (caller workspace)
        got = torch.full_like(want, float("nan"))                   # caller-owned output, allocated before the call
        torch.cuda.synchronize()
        before = torch.cuda.memory_allocated()
        rc = _static_forward(se_engine, partials, got)
        assert rc == L.AVC_OK, L.last_error(se_engine._h)
        assert torch.cuda.memory_allocated() == before              # nothing was taken from torch's allocator either
        torch.cuda.synchronize()
        err = float((got - want).abs().max())
        print(f"call {i}: max|static - workspace| = {err:.2e}")
        assert torch.isfinite(got).all() and err <= BF16_TOL
        assert float((got.norm(dim=1) - 1).abs().max()) <= 1e-5
    # the same input again after other shapes have used the array: identical result
    again = torch.empty(8, 256, device="cuda")
    assert _static_forward(se_engine, batches[0], again) == L.AVC_OK
    torch.cuda.synchronize()
    assert float((again - se_engine.se_forward(batches[0])).abs().max()) <= BF16_TOL


def test_static_scratch_entry_point_rejects_sizes_beyond_the_array(se_engine):
    from autovc_b200 import _lib as L
    partials = torch.zeros(16, 160, 40, device="cuda")              # needs 17.8 MiB > 12 MiB
    out = torch.empty(16, 256, device="cuda")
    assert int(L.lib().avc_se_workspace_bytes(se_engine._h, 16, 160)) > (12 << 20)
    rc = _static_forward(se_engine, partials, out)
    assert rc == L.AVC_ERR_WORKSPACE
    msg = L.last_error(se_engine._h)
    assert "static" in msg and "avc_se_forward" in msg and str(12 << 20) in msg
    assert _static_forward(se_engine, partials[:0].reshape(0, 160, 40), out) == L.AVC_ERR_INVALID_ARG
