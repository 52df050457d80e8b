"""GPU probe of the batch-statistics forward against the reference goldens: error per output and per running buffer."""
import os, sys
import numpy as np
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import autovc_b200
from autovc_b200.synth import synth_inputs, synth_state_dict
from conftest import load_golden
P = [f"encoder.convolutions.{i}" for i in range(3)] + [f"decoder.convolutions.{i}" for i in range(3)] + [f"postnet.convolutions.{i}" for i in range(5)]
for name in ("bn_train_b4_t64", "bn_train_b3_t100_w3"):
    g = load_golden(name)
    B, T = int(g["B"]), int(g["T"])
    for prec in ("fp32", "bf16"):
        m = autovc_b200.AutoEncoder(device="cuda", verbose=False, precision=prec, batch_stat_bn=True)
        m.load_state_dict(synth_state_dict(int(g["wseed"]))); m.train()
        x, c_org, c_trg = (t.cuda() for t in synth_inputs(B, T, seed=int(g["xseed"])))
        mel, post, codes = m(x, c_org, c_trg)
        err = {k: float((v.cpu() - torch.from_numpy(g[n])).abs().max()) for k, v, n in (("mel", mel, "mel"), ("post", post, "mel_post"), ("codes", codes, "codes"))}
        cos = float(torch.nn.functional.cosine_similarity(codes.cpu(), torch.from_numpy(g["codes"]), dim=1).min())
        sd = m.state_dict()
        rm = [float(np.abs(sd[p + ".1.running_mean"].numpy() - g[f"p1.rm{i}"]).max()) for i, p in enumerate(P)]
        rv = [float(np.abs(sd[p + ".1.running_var"].numpy() / g[f"p1.rv{i}"] - 1).max()) for i, p in enumerate(P)]
        nbt = [int(sd[p + ".1.num_batches_tracked"]) - int(g[f"p1.nbt{i}"]) for i, p in enumerate(P)]
        print(name, prec, {k: f"{v:.2e}" for k, v in err.items()}, f"cos {cos:.6f}")
        print("   rm err", " ".join(f"{v:.1e}" for v in rm)); print("   rv rel", " ".join(f"{v:.1e}" for v in rv)); print("   nbt diff", nbt)
        c2 = m(torch.from_numpy(g["mel_post"]).cuda(), c_org, None)
        print("   pass2 codes err %.2e cos %.6f" % (float((c2.cpu() - torch.from_numpy(g["codes_pass2"])).abs().max()),
              float(torch.nn.functional.cosine_similarity(c2.cpu(), torch.from_numpy(g["codes_pass2"]), dim=1).min())))
