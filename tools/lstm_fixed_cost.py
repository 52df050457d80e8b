"""Fixed cost of a recurrence launch: phase times at several T (B=256): time = fixed + (T-1) * step."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, autovc_b200
from autovc_b200.synth import synth_inputs, synth_state_dict
dev = torch.device("cuda", 0)
m = autovc_b200.AutoEncoder(device=dev, verbose=False); m.load_state_dict(synth_state_dict(0))
eng = None
res = {}
for T in (32, 64, 128, 256):
    x, c1, c2 = (t.to(dev) for t in synth_inputs(256, T, seed=1))
    for _ in range(3): m(x, c1, c2)
    eng = m._engine(dev); eng.set_profiling(True)
    acc = {}
    for _ in range(8):
        m(x, c1, c2)
        for k, v in eng.profile().items(): acc[k] = acc.get(k, 0) + v / 8
    eng.set_profiling(False)
    res[T] = acc
    print(T, {k: round(v, 4) for k, v in acc.items()}, flush=True)
for ph, nl in (("lstm2", 2), ("lstm1", 1)):
    a, b = res[32][ph], res[256][ph]
    step = (b - a) / (256 - 32) / nl
    print(f"{ph}: step {step * 1e3:.3f} us, fixed per launch {(a / nl - 31 * step) * 1e3:.1f} us")
