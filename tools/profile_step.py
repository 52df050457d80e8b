"""Minimal driver for ncu: W warm-up forwards then K forwards of config 2 (B=256, T=128, bf16)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import autovc_b200
from autovc_b200.synth import synth_inputs, synth_state_dict

B = int(os.environ.get("PB", 256)); T = int(os.environ.get("PT", 128))
W = int(os.environ.get("PW", 2)); K = int(os.environ.get("PK", 1))
m = autovc_b200.AutoEncoder(device="cuda", verbose=False)
m.load_state_dict(synth_state_dict(0))
x, c1, c2 = (t.cuda() for t in synth_inputs(B, T))
for _ in range(W + K):
    out = m(x, c1, c2)
torch.cuda.synchronize()
print("launches per forward:", m.last_launch_count, float(out[1].abs().max()))
