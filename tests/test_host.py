"""Host-side logic that needs no GPU: the C-ABI library loads and exports every declared symbol, weight
folding/packing (done in C++ inside avc_load_weights, inspected through a host-only context), the module's
state-dict contract, error behaviour."""
import ctypes
import os
import re

import numpy as np
import pytest
import torch

from autovc_b200.synth import state_dict_spec, synth_state_dict
from oracle import autovc_oracle as O

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_every_declared_symbol(built_lib):
    header = open(os.path.join(ROOT, "include", "autovc_b200.h")).read()
    declared = sorted(set(re.findall(r"\b(avc_[a-z_]+)\s*\(", header)))
    assert len(declared) >= 10
    lib = ctypes.CDLL(built_lib)
    for name in declared:
        assert hasattr(lib, name), f"{name} declared in include/autovc_b200.h but not exported"
    assert lib.avc_abi_version() == 8


def test_state_dict_contract():
    import autovc_b200
    m = autovc_b200.AutoEncoder(device="cpu", verbose=False)
    keys = list(m.state_dict().keys())
    assert keys == [n for n, _, _ in state_dict_spec()]
    for (n, shape, _), (k, v) in zip(state_dict_spec(), m.state_dict().items()):
        assert tuple(v.shape) == tuple(shape), n
    assert sum(p.numel() for p in m.parameters()) == 28_422_464
    assert m.encoder.freq == 32 and m.decoder.linear_projection.linear_layer.out_features == 80
    assert not m.training


@pytest.fixture(scope="module")
def host_engine(built_lib):
    from autovc_b200.engine import Engine
    eng = Engine((32, 256, 512, 32), -1, "fp32")   # device -1: host-only context (pack, no run)
    eng.load_state_dict(synth_state_dict(0))
    yield eng
    eng.close()


def test_conv_bn_fold_and_tap_major_packing(host_engine):
    sd = synth_state_dict(0)
    p = "encoder.convolutions.0"
    w, b = O.fold_batchnorm(sd[p + ".0.conv.weight"].double(), sd[p + ".0.conv.bias"].double(), sd[p + ".1.weight"].double(),
                            sd[p + ".1.bias"].double(), sd[p + ".1.running_mean"].double(), sd[p + ".1.running_var"].double())
    packed = host_engine.packed("enc.conv0.w").view(512, 5, 384)       # [Cout][tap][Cin padded to 384]
    assert torch.allclose(packed[:, :, :336], w.permute(0, 2, 1).float(), atol=1e-7)
    assert float(packed[:, :, 336:].abs().max()) == 0.0
    assert torch.allclose(host_engine.packed("enc.conv0.b"), b.float(), atol=1e-7)
    # postnet first layer: Cin 80 padded to 128
    assert host_engine.packed("post.conv0.w").numel() == 512 * 5 * 128
    assert host_engine.packed("post.conv4.w").numel() == 80 * 5 * 512


def test_unfolded_conv_packing_for_batch_statistics_mode(host_engine):
    """The train-mode BatchNorm forward (avc_forward_bn_train) runs the convolutions WITHOUT the fold: same tap-major
    layout, the conv's own weights and bias."""
    sd = synth_state_dict(0)
    for blob, p, cin, cp, cout in (("raw.enc.conv0", "encoder.convolutions.0", 336, 384, 512),
                                   ("raw.dec.conv2", "decoder.convolutions.2", 512, 512, 512),
                                   ("raw.post.conv4", "postnet.convolutions.4", 512, 512, 80)):
        packed = host_engine.packed(blob + ".w").view(cout, 5, cp)
        assert torch.equal(packed[:, :, :cin], sd[p + ".0.conv.weight"].permute(0, 2, 1))
        assert torch.equal(host_engine.packed(blob + ".b"), sd[p + ".0.conv.bias"])
    assert not torch.equal(host_engine.packed("raw.enc.conv0.w"), host_engine.packed("enc.conv0.w"))


def test_lstm_unit_major_gate_packing(host_engine):
    sd = synth_state_dict(0)
    H = 1024
    whh = sd["decoder.lstm2.weight_hh_l1"]
    packed = host_engine.packed("dec.lstm2.l1.hh.w").view(H, 4, H)     # row u*4+g <- torch row g*H+u
    for g in range(4):
        assert torch.equal(packed[:, g, :], whh[g * H:(g + 1) * H, :])
    bias = host_engine.packed("dec.lstm2.l1.ih.b").view(H, 4)
    full = sd["decoder.lstm2.bias_ih_l1"] + sd["decoder.lstm2.bias_hh_l1"]
    assert torch.equal(bias[:, 2], full[2 * H:3 * H])
    # bidirectional encoder: direction-major, then unit-major
    wih = host_engine.packed("enc.lstm.l1.ih.w").view(2, 32, 4, 64)
    assert torch.equal(wih[1, :, 3, :], sd["encoder.lstm.weight_ih_l1_reverse"][96:128, :])


def test_missing_weight_is_reported(built_lib):
    from autovc_b200.engine import Engine
    eng = Engine((32, 256, 512, 32), -1, "bf16")
    sd = synth_state_dict(0)
    del sd["decoder.lstm1.weight_hh_l0"]
    with pytest.raises(KeyError, match="decoder.lstm1.weight_hh_l0"):
        eng.load_state_dict(sd)
    eng.close()


def test_unsupported_dims_and_host_only_run(built_lib):
    from autovc_b200.engine import Engine
    with pytest.raises(RuntimeError, match="dim_neck"):
        Engine((30, 256, 512, 32), -1, "bf16")
    eng = Engine((32, 256, 512, 32), -1, "bf16")
    assert eng.workspace_bytes(256, 128) > 0
    eng.close()


def test_no_cpu_fallback():
    import autovc_b200
    m = autovc_b200.AutoEncoder(device="cpu", verbose=False)
    with pytest.raises(RuntimeError, match="CUDA"):
        m(torch.zeros(1, 80, 64), torch.zeros(1, 256), torch.zeros(1, 256))
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        torch.ops.autovc_b200.generator_encode(torch.zeros(1, 80, 64), torch.zeros(1, 256), 1)


def test_overlap_average_matches_oracle():
    import autovc_b200
    x = torch.randn(4, 80, 64)
    assert torch.allclose(autovc_b200.overlap_average(x, 0.5), O.overlap_average(x, 0.5))
    assert autovc_b200.overlap_average(x, 0.25).shape == (80, 64 + 3 * 48)


def test_vocoder_state_dict_contract_and_host_packing(built_lib):
    """The vocoder module carries the reference's state-dict names (strict load of the seeded dict) and the C ABI folds
    and packs them on a host-only handle (no GPU): a missing entry is an error that names it."""
    from autovc_b200 import _lib as L
    from autovc_b200.synth import synth_wavernn_state_dict
    from autovc_b200.vocoder import WaveRNN, fold_params
    sd = synth_wavernn_state_dict(11)
    m = WaveRNN(device="cpu")
    res = m.load_state_dict(sd, strict=True)
    assert not res.missing_keys and not res.unexpected_keys
    assert m.state_dict()["upsample.up_layers.5.weight"].shape == (1, 1, 1, 23)
    h = ctypes.c_void_p()
    assert L.lib().avc_voc_create(ctypes.byref(h), -1) == L.AVC_OK
    keep = {k: v.float().contiguous() for k, v in sd.items() if k != "step" and not k.endswith("num_batches_tracked")}

    def load(d):
        table = (L.AvcWeightDesc * len(d))()
        for i, (k, v) in enumerate(d.items()):
            table[i] = L.AvcWeightDesc(k.encode(), ctypes.cast(v.data_ptr(), ctypes.POINTER(ctypes.c_float)), v.numel())
        return L.lib().avc_voc_load_weights(h, table, len(d))

    assert load(keep) == L.AVC_OK
    bad = dict(keep)
    del bad["rnn2.weight_hh_l0"]
    assert load(bad) == L.AVC_ERR_MISSING_WEIGHT and b"rnn2.weight_hh_l0" in L.lib().avc_voc_last_error(h)
    bad = dict(keep)
    bad["I.weight"] = bad["I.weight"][:, :112].contiguous()
    assert load(bad) == L.AVC_ERR_MISSING_WEIGHT and b"I.weight" in L.lib().avc_voc_last_error(h)
    L.lib().avc_voc_destroy(h)
    # the fold arithmetic of generate (model.py:381-391)
    assert fold_params(35200, 11000, 550) == (3, 0) and fold_params(6050, 2950, 50) == (2, 0) and fold_params(6600, 1500, 100) == (5, 1600)
    with pytest.raises(RuntimeError):
        WaveRNN(device="cpu").condition(torch.zeros(1, 80, 30))


def test_weight_tracking_key_follows_in_place_updates_and_invalidate():
    """PackedWeightsMixin: the key the modules compare before re-packing their weights (no GPU needed)."""
    import torch
    import autovc_b200
    m = autovc_b200.AutoEncoder(device="cpu", verbose=False)
    k0 = m._weights_key()
    p = next(m.parameters())
    p.data.mul_(1.0)                                   # a write through .data is invisible to Tensor._version ...
    assert m._weights_key() == k0
    m.invalidate()                                     # ... the documented hook covers it
    k1 = m._weights_key()
    assert k1 != k0
    with torch.no_grad():
        p.mul_(1.0)
    k2 = m._weights_key()
    assert k2 != k1
    m.encoder.convolutions[0][1].running_mean.add_(0.0)
    k3 = m._weights_key()
    assert k3[0] != k2[0] and k3[1] == k2[1]           # running statistics only touch the first key
    flat = m.flatten_params()
    assert flat.numel() == sum(q.numel() for q in m.parameters())
    m.load_flattened_params(flat.clone())
    assert m._weights_key() != k3
    m.load_state_dict(m.state_dict())
    assert m._weights_key()[0][0] > k3[0][0]
    for cls in (autovc_b200.SpeakerEncoder, autovc_b200.WaveRNN):
        mod = cls(device="cpu") if cls is autovc_b200.WaveRNN else cls(device="cpu", verbose=False)
        a = mod._weights_key()
        mod.invalidate()
        assert mod._weights_key() != a


def test_staged_reference_matches_its_manifest():
    """oracle/_ref (staged by oracle/make_ref.py in the build container) is the unmodified reference: every file's
    SHA-256 still equals the one recorded at copy time, and - where /root/reference is present - the source's."""
    import hashlib
    import json
    import os
    import pytest
    from oracle import ref_loader
    if not os.path.isdir(os.path.join(ref_loader.STAGED, "autovc")):
        pytest.skip("oracle/_ref not staged (run python oracle/make_ref.py in the build container)")
    assert ref_loader.verify()
    files = json.load(open(os.path.join(ref_loader.STAGED, "MANIFEST.json")))["files"]
    assert "autovc/auto_encoder/model.py" in files
    if os.path.isdir(ref_loader.LIVE):
        for rel, sha in files.items():
            with open(os.path.join(ref_loader.LIVE, rel), "rb") as f:
                assert hashlib.sha256(f.read()).hexdigest() == sha, rel


def test_conv0_split_for_fused_speaker_concat(host_engine):
    """enc.conv0 split along the input channels (reference encoder.py:65-66 fused away): the mel part stays a 5-tap
    convolution over 80 (padded to 128) channels, the speaker part becomes a [5*512][256] matrix (row tap*512 + n)."""
    full = host_engine.packed("enc.conv0.w").view(512, 5, 384)
    mel = host_engine.packed("enc.conv0.mel.w").view(512, 5, 128)
    spk = host_engine.packed("enc.conv0.spk.w").view(5, 512, 256)
    assert torch.equal(mel[:, :, :80], full[:, :, :80]) and float(mel[:, :, 80:].abs().max()) == 0.0
    assert torch.equal(spk.permute(1, 0, 2), full[:, :, 80:336])
    assert torch.equal(host_engine.packed("enc.conv0.mel.b"), host_engine.packed("enc.conv0.b"))
    assert float(host_engine.packed("enc.conv0.spk.b").abs().max()) == 0.0
