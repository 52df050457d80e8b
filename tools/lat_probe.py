"""Phase times of a single 128-frame utterance (B=1): where the p50 latency goes."""
import os, sys, statistics
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, autovc_b200
from autovc_b200.synth import synth_inputs, synth_state_dict
B = int(os.environ.get("PB", 1)); T = int(os.environ.get("PT", 128))
m = autovc_b200.AutoEncoder(device="cuda", verbose=False); m.load_state_dict(synth_state_dict(0))
x, c1, c2 = (t.cuda() for t in synth_inputs(B, T))
for _ in range(5): m(x, c1, c2)
eng = m._engine(x.device)
eng.set_profiling(True)
acc = {}
for _ in range(10):
    m(x, c1, c2)
    for k, v in eng.profile().items(): acc[k] = acc.get(k, 0) + v / 10
eng.set_profiling(False)
lat = []
for _ in range(50):
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record(); m(x, c1, c2); e.record(); e.synchronize(); lat.append(s.elapsed_time(e))
print(f"B={B} T={T}: p50 {statistics.median(lat):.3f} ms; phases (ms):", {k: round(v, 4) for k, v in acc.items()}, "launches", m.last_launch_count)
