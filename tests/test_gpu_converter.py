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


def test_convert_batch_pipelined_from_host_rows():
    """Host-resident mels go through HostPipeline (staged or, for consecutive rows of one pinned tensor, copied straight from
    it) and give exactly what plain forward calls give, on the device or back on the host, in input order across
    length buckets."""
    import autovc_b200
    from autovc_b200.voice_converter import host_rows
    m = autovc_b200.AutoEncoder(device="cuda", verbose=False, precision="bf16")
    m.load_state_dict(synth_state_dict(0))
    vc = VoiceConverter(m, device="cuda", max_batch=8)
    xa, ca, ta = synth_inputs(20, 64, seed=300)        # 3 batches of a 64-frame bucket (8, 8, 4)
    xb, cb, tb = synth_inputs(5, 100, seed=301)        # one ragged bucket
    slab = xa.pin_memory()
    assert host_rows([slab[i] for i in range(4, 12)]).data_ptr() == slab[4].data_ptr()          # zero-copy view
    assert host_rows([slab[i] for i in (1, 3)]).data_ptr() != slab[1].data_ptr()                # staged copy
    mels = [slab[i] for i in range(20)] + [xb[i] for i in range(5)]
    cs = [ca[i] for i in range(20)] + [cb[i] for i in range(5)]
    ct = [ta[i] for i in range(20)] + [tb[i] for i in range(5)]
    order = [21, 0, 1, 2, 3, 22, 4, 5, 6, 7, 8, 9, 10, 11, 20, 12, 13, 14, 15, 16, 17, 18, 19, 23, 24]
    mels, cs, ct = [mels[i] for i in order], [cs[i] for i in order], [ct[i] for i in order]
    want = [m(x[None].cuda(), a[None].cuda(), b[None].cuda())[1][0].cpu() for x, a, b in zip(mels, cs, ct)]
    for output in ("device", "host"):
        got = vc.convert_batch(mels, cs, ct, output=output)
        torch.cuda.synchronize()
        assert all(g.is_cuda == (output == "device") for g in got)
        for g, w in zip(got, want):
            assert float((g.cpu() - w).abs().max()) <= 1e-5


def test_learn_speakers_embeds_each_speaker_once_in_one_batch():
    import autovc_b200
    from autovc_b200.synth import synth_speaker_encoder
    se_sd, parts = synth_speaker_encoder(5)
    se = autovc_b200.SpeakerEncoder(device="cuda", verbose=False)
    se.load_state_dict(se_sd)
    vc = VoiceConverter(None, embed_speaker=se.as_embed_fn(), device="cuda")
    datas = [parts * (1.0 + 0.1 * k) for k in range(4)]
    keys = ["a", "b", "a", "c", "b", "d"]
    n = vc.learn_speakers(keys, [datas["abcd".index(k)] for k in keys], embed_many=se.embed_speakers)
    assert n == 4 and vc.embed_calls == 4 and set(vc.speakers) == set("abcd")
    for k in "abcd":
        one = se.embed_partials(datas["abcd".index(k)])
        assert float((vc.speakers[k] - one).abs().max()) <= 2e-3
    assert vc.learn_speakers(keys, [None] * 6) == 0
