"""Host logic of the batched conversion (no GPU): pairing order, once-per-speaker embedding, length bucketing without
padding, chunk + overlap-average, contiguous sharding, and the world-size-2 gloo path.  The generator itself is replaced
by the CPU oracle here (test infrastructure), so the results can be compared pair by pair."""
import os

import pytest
import torch
import torch.multiprocessing as mp

from autovc_b200.synth import synth_inputs, synth_state_dict
from autovc_b200.voice_converter import Utterance, VoiceConverter, bucket_by_length, chunk_mel, shard_range
from oracle import autovc_oracle as O


class OracleAE:
    """Same forward contract as autovc_b200.AutoEncoder, computed by the oracle; counts calls and batch sizes."""
    device = torch.device("cpu")

    def __init__(self):
        self.sd = synth_state_dict(0)
        self.calls = []

    def __call__(self, x, c_org, c_trg):
        self.calls.append(tuple(x.shape))
        return O.generator_forward(self.sd, x, c_org, c_trg)


def make_world(n_src=3, n_trg=2, lengths=(64, 96, 64)):
    srcs = []
    for i in range(n_src):
        x, _, _ = synth_inputs(1, lengths[i % len(lengths)], seed=100 + i)
        _, c, _ = synth_inputs(1, 32, seed=300 + i % 2)                     # speaker data belongs to the speaker
        srcs.append(Utterance(x[0], f"src{i % 2}", speaker_data=c[0]))      # two source speakers for three utterances
    trgs = []
    for j in range(n_trg):
        x, c, _ = synth_inputs(1, 64, seed=200 + j)
        trgs.append(Utterance(x[0], f"trg{j}", speaker_data=c[0]))
    return srcs, trgs


def test_shard_and_bucket_helpers():
    assert [list(shard_range(10, r, 4)) for r in range(4)] == [[0, 1, 2], [3, 4, 5], [6, 7, 8], [9]]
    assert list(shard_range(2, 3, 4)) == []
    assert bucket_by_length([64, 96, 64, 32]) == {64: [0, 2], 96: [1], 32: [3]}
    assert chunk_mel(torch.zeros(80, 400), 128, 0.5).shape == (5, 80, 128)
    with pytest.raises(ValueError):
        chunk_mel(torch.zeros(80, 100), 128)


def test_convert_multiple_order_dedupe_and_bucketing():
    ae = OracleAE()
    vc = VoiceConverter(ae, embed_speaker=lambda d: d, max_batch=4)
    srcs, trgs = make_world()
    out = vc.convert_multiple(srcs, trgs, "all_combinations")
    assert len(out) == 6
    # every speaker embedded exactly once (2 source speakers + 2 targets), not once per pair like the reference loop
    assert vc.embed_calls == 4
    # one forward per length bucket (64: 4 pairs, 96: 2 pairs), never a padded batch
    assert sorted(ae.calls) == [(2, 80, 96), (4, 80, 64)]
    # reference order: product(sources, targets); each result equals the single-pair conversion
    k = 0
    for s in srcs:
        for t in trgs:
            _, want, _ = O.generator_forward(ae.sd, s.mel[None], vc.speakers[s.speaker][None], vc.speakers[t.speaker][None])
            assert out[k].shape == s.mel.shape
            assert float((out[k] - want[0]).abs().max()) < 1e-5
            k += 1


def test_align_bidirectional_and_known_key_targets():
    ae = OracleAE()
    vc = VoiceConverter(ae, embed_speaker=lambda d: d)
    srcs, trgs = make_world(n_src=2, n_trg=2, lengths=(64,))
    out = vc.convert_multiple(srcs, trgs, "align", bidirectional=True)
    assert len(out) == 4                                     # 2 forward pairs + 2 reversed
    _, want, _ = O.generator_forward(ae.sd, trgs[1].mel[None], trgs[1].speaker_data[None], srcs[1].speaker_data[None])
    assert float((out[3] - want[0]).abs().max()) < 1e-5
    vc.learn_speaker("mean_speaker", trgs[0].speaker_data)
    out2 = vc.convert_multiple(srcs, ["mean_speaker"])
    assert float((out2[0] - out[0]).abs().max()) < 1e-5   # batch of 2 vs batch of 4: BLAS blocking differs
    with pytest.raises(AssertionError):
        vc.convert_multiple(srcs, ["mean_speaker"], bidirectional=True)   # reference `:376`
    with pytest.raises(KeyError):
        vc.convert(srcs[0], "nobody")


def test_convert_cut_matches_batch_forward_semantics():
    ae = OracleAE()
    vc = VoiceConverter(ae, embed_speaker=lambda d: d)
    x, c1, c2 = synth_inputs(1, 256, seed=9)
    u = Utterance(x[0], "a", c1[0])
    got = vc.convert(u, c2[0], cut=True, n_frames=128, overlap=0.5)
    want = O.batch_forward(ae.sd, chunk_mel(x[0], 128, 0.5), c1, c2, overlap=0.5)
    assert got.shape == (80, 256) and float((got - want).abs().max()) < 1e-5


def _dist_worker(rank, world, port, q):
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    torch.set_num_threads(2)
    ae = OracleAE()
    vc = VoiceConverter(ae, embed_speaker=lambda d: d)
    srcs, trgs = make_world()
    out = vc.convert_multiple(srcs, trgs, distributed=True, gather=True)
    n_local = sum(c[0] for c in ae.calls)
    q.put((rank, n_local, [o.sum().item() for o in out]))
    dist.barrier()
    dist.destroy_process_group()


def test_world_size_2_sharding_gloo():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + os.getpid() % 1000
    procs = [ctx.Process(target=_dist_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = sorted(q.get(timeout=240) for _ in range(2))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    # each rank converted its contiguous half (3 of 6 pairs), and after the gather both hold the same full list
    assert got[0][1] == 3 and got[1][1] == 3
    assert got[0][2] == pytest.approx(got[1][2])
    ae = OracleAE()
    ref = VoiceConverter(ae, embed_speaker=lambda d: d).convert_multiple(*make_world())
    assert [o.sum().item() for o in ref] == pytest.approx(got[0][2], rel=1e-4)
