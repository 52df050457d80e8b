"""Generate `tests/golden/bn_train_*.npz` by running the UNMODIFIED reference module in TRAINING mode (BatchNorm on
batch statistics), i.e. in the state `AutoEncoder.learn` runs it (`/root/reference/autovc/auto_encoder/model.py:306`
forward, `:201` content-code re-encode of the postnet output) and `VoiceConverter.convert` leaves it.
Build-container only:  python oracle/make_golden_bn_train.py

Stored per case: the three outputs of pass 1 (`forward(x, c, c_trg)`), the codes of pass 2 (`forward(mel_post, c, None)`,
also in training mode), and every BatchNorm running buffer after pass 1 and after pass 2 (11 layers; pass 2 touches the
encoder's three).  Weights and inputs are regenerated from the seeds by the tests.
"""
from __future__ import annotations

import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))
from make_golden import import_reference, OUT  # noqa: E402

BN_PREFIXES = [f"encoder.convolutions.{i}" for i in range(3)] + [f"decoder.convolutions.{i}" for i in range(3)] + \
              [f"postnet.convolutions.{i}" for i in range(5)]


def running(model):
    sd = model.state_dict()
    out = {}
    for i, p in enumerate(BN_PREFIXES):
        out[f"rm{i}"] = sd[p + ".1.running_mean"].numpy().copy()
        out[f"rv{i}"] = sd[p + ".1.running_var"].numpy().copy()
        out[f"nbt{i}"] = int(sd[p + ".1.num_batches_tracked"])
    return out


def main():
    from autovc_b200.synth import synth_state_dict, synth_inputs
    AutoEncoder, _ = import_reference()
    torch.set_num_threads(8)
    for name, B, T, wseed, xseed in (("bn_train_b4_t64", 4, 64, 0, 31), ("bn_train_b3_t100_w3", 3, 100, 3, 32)):
        sd = synth_state_dict(wseed)
        model = AutoEncoder(device="cpu", verbose=False)
        model.load_state_dict(sd)
        model.train()
        x, c_org, c_trg = synth_inputs(B, T, seed=xseed)
        with torch.no_grad():
            mel, post, codes = model(x, c_org, c_trg)
            r1 = running(model)
            codes2 = model(post, c_org, None)
            r2 = running(model)
        out = dict(B=B, T=T, wseed=wseed, xseed=xseed, mel=mel.numpy(), mel_post=post.numpy(), codes=codes.numpy(),
                   codes_pass2=codes2.numpy())
        out.update({"p1." + k: v for k, v in r1.items()})
        out.update({"p2." + k: v for k, v in r2.items()})
        np.savez_compressed(os.path.join(OUT, name + ".npz"), **out)
        print(name, "post|max|=%.4f codes|max|=%.4f nbt(enc0)=%d" % (post.abs().max(), codes.abs().max(), r2["nbt0"]))


if __name__ == "__main__":
    main()
