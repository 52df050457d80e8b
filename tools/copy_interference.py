"""N=1: does a DMA copy in flight slow the generator's kernels?  Step time + phase times with no copy, with a continuous D2H
loop and with a continuous H2D loop on a side stream."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, autovc_b200
from autovc_b200.synth import synth_inputs, synth_state_dict
dev = torch.device("cuda", 0)
m = autovc_b200.AutoEncoder(device=dev, verbose=False); m.load_state_dict(synth_state_dict(0))
x, c1, c2 = (t.to(dev) for t in synth_inputs(256, 128, seed=1))
hbuf = torch.empty(21 << 20, dtype=torch.uint8).pin_memory(); dbuf = torch.empty(21 << 20, dtype=torch.uint8, device=dev)
side = torch.cuda.Stream()
eng = m._engine(dev)
for _ in range(3): m(x, c1, c2)
def run(mode, n_copy):
    torch.cuda.synchronize()
    eng.set_profiling(True)
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    acc = {}
    tot = 0.0
    for it in range(10):
        with torch.cuda.stream(side):
            for _ in range(n_copy):
                if mode == "d2h": hbuf.copy_(dbuf, non_blocking=True)
                elif mode == "h2d": dbuf.copy_(hbuf, non_blocking=True)
        s.record(); m(x, c1, c2); e.record(); e.synchronize(); side.synchronize()
        tot += s.elapsed_time(e)
        for k, v in eng.profile().items(): acc[k] = acc.get(k, 0) + v / 10
    eng.set_profiling(False)
    print(f"{mode:5s} x{n_copy}: step {tot / 10:.3f} ms  " + " ".join(f"{k} {v:.3f}" for k, v in acc.items()), flush=True)
run("none", 0); run("d2h", 1); run("d2h", 6); run("h2d", 11); run("none", 0)

# ---- the same with the forward replayed as ONE CUDA graph ----
out = tuple(torch.empty_like(t) for t in m(x, c1, c2))
torch.cuda.synchronize()
try:
    g = torch.cuda.CUDAGraph()
    cs = torch.cuda.Stream()
    with torch.cuda.stream(cs):
        m(x, c1, c2, out=out)
        torch.cuda.synchronize()
        with torch.cuda.graph(g, stream=cs):
            m(x, c1, c2, out=out)
    torch.cuda.synchronize()
    ref = [t.clone() for t in m(x, c1, c2)]
    def run_g(mode, n_copy):
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        tot = 0.0
        for it in range(10):
            with torch.cuda.stream(side):
                for _ in range(n_copy):
                    if mode == "d2h": hbuf.copy_(dbuf, non_blocking=True)
                    elif mode == "h2d": dbuf.copy_(hbuf, non_blocking=True)
            s.record(); g.replay(); e.record(); e.synchronize(); side.synchronize()
            tot += s.elapsed_time(e)
        print(f"graph {mode:5s} x{n_copy}: step {tot / 10:.3f} ms  equal {all(torch.equal(a, b) for a, b in zip(out, ref))}", flush=True)
    run_g("none", 0); run_g("d2h", 1); run_g("d2h", 6); run_g("h2d", 11)
except Exception as exc:
    print("graph capture failed:", type(exc).__name__, str(exc)[:300])
