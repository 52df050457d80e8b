"""One-off GPU fuzz: for random (B, T) the bf16 results of rows inside a batch must equal, bit for bit, the results of the same
rows run alone and in a small batch (different kernel variants: one / two sub-tiles, 8 / 16 group warps, N = 16 / 32 MMAs, ragged
tiles).   python tools/fuzz_rows.py [cases] [seed]"""
import random
import sys

import torch

from autovc_b200 import AutoEncoder
from autovc_b200.synth import synth_inputs, synth_state_dict

cases = int(sys.argv[1]) if len(sys.argv) > 1 else 30
rng = random.Random(int(sys.argv[2]) if len(sys.argv) > 2 else 0)
m = AutoEncoder(verbose=False, device="cuda:0", precision="bf16")
m.load_state_dict(synth_state_dict(0))
m.eval()
bad = 0
for case in range(cases):
    B = rng.choice([1, 2, 5, 8, 9, 16, 17, 31, 32, 33, 40, 63, 64, 65, 96, 128, 129, 160, 200, 256, 257, 300])
    T = rng.choice([32, 33, 45, 64, 96, 100, 128, 129, 200, 257])
    x, c_org, c_trg = (t.cuda() for t in synth_inputs(B, T, seed=5000 + case))
    full = [t.clone() for t in m(x, c_org, c_trg)]
    rows = sorted({0, B - 1, rng.randrange(B), rng.randrange(B)})
    ok = True
    for r in rows:
        one = m(x[r:r + 1], c_org[r:r + 1], c_trg[r:r + 1])
        ok &= all(torch.equal(a[0], f[r]) for a, f in zip(one, full))
    k = min(B, 9)
    sub = m(x[:k], c_org[:k], c_trg[:k])
    ok &= all(torch.equal(a, f[:k]) for a, f in zip(sub, full))
    ok &= bool(torch.isfinite(full[1]).all())
    print(f"case {case}: B={B} T={T} rows={rows} {'ok' if ok else 'MISMATCH'}", flush=True)
    bad += not ok
print("mismatches:", bad)
sys.exit(1 if bad else 0)
