"""GPU tests of the batched conversion layer: VoiceConverter on the CUDA generator against the CPU oracle, and the
HostPipeline (copy/compute overlap) against plain forward calls."""
import pytest
import torch

from autovc_b200.synth import synth_inputs, synth_state_dict
from autovc_b200.voice_converter import HostPipeline, Utterance, VoiceConverter

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ae():
    import autovc_b200
    m = autovc_b200.AutoEncoder(device="cuda", verbose=False, precision="fp32")
    m.load_state_dict(synth_state_dict(0))
    return m


def test_convert_multiple_matches_oracle_pairwise(ae):
    from oracle import autovc_oracle as O
    sd = synth_state_dict(0)
    vc = VoiceConverter(ae, embed_speaker=lambda d: d, device="cuda", max_batch=8)
    srcs, trgs = [], []
    for i, T in enumerate((64, 100, 64, 33)):
        x, _, _ = synth_inputs(1, T, seed=40 + i)
        _, c, _ = synth_inputs(1, 32, seed=60 + i % 2)
        srcs.append(Utterance(x[0], f"s{i % 2}", c[0]))
    for j in range(3):
        _, c, _ = synth_inputs(1, 32, seed=80 + j)
        trgs.append(Utterance(torch.zeros(80, 32), f"t{j}", c[0]))
    out = vc.convert_multiple(srcs, trgs)
    assert len(out) == 12 and vc.embed_calls == 5
    k = 0
    for s in srcs:
        for t in trgs:
            _, want, _ = O.generator_forward(sd, s.mel[None], vc.speakers[s.speaker][None], vc.speakers[t.speaker][None])
            assert float((out[k].cpu() - want[0]).abs().max()) <= 1e-4
            k += 1
    got = vc.convert(srcs[1], "t0")
    assert float((got - out[3]).abs().max()) <= 1e-5


def test_host_pipeline_equals_plain_forward():
    import autovc_b200
    m = autovc_b200.AutoEncoder(device="cuda", verbose=False, precision="bf16")
    m.load_state_dict(synth_state_dict(0))
    dev = torch.device("cuda", 0)
    batches, outs, want = [], [], []
    for i in range(5):
        x, c1, c2 = synth_inputs(8, 128, seed=500 + i)
        batches.append((x.pin_memory(), c1.pin_memory(), c2.pin_memory()))
        outs.append((torch.empty(8, 80, 128).pin_memory(), torch.empty(8, 80, 128).pin_memory(), torch.empty(8, 256).pin_memory()))
        want.append(tuple(t.cpu() for t in m(x.cuda(), c1.cuda(), c2.cuda())))
    HostPipeline(m, dev).run(batches, outs)
    torch.cuda.synchronize()
    for o, w in zip(outs, want):
        for a, b in zip(o, w):
            assert torch.equal(a, b)
