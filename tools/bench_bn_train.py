"""Training-shape forward pair (BASELINE config 5's per-GPU share: B=256, T=128) with BatchNorm on batch statistics
 next to the eval-mode pair.  One JSON line.   python tools/bench_bn_train.py [--B 256 --T 128]"""
import argparse, json, os, sys
import numpy as np
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import autovc_b200
from autovc_b200.synth import synth_inputs, synth_state_dict

FLOP_PAIR = 64_061_440   # full forward + encoder-only pass, per frame


def timed(fn, n=20, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(n):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    return float(np.median(ts))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--B", type=int, default=256)
    ap.add_argument("--T", type=int, default=128)
    a = ap.parse_args()
    sd = synth_state_dict(0)
    x, c, _ = (t.cuda() for t in synth_inputs(a.B, a.T, seed=3))
    out = {"metric": "training-shape forward pair, frames/s", "config": {"workload": f"B={a.B} x T={a.T}: forward(x,c,c) + forward(mel_post,c,None)"}}
    for mode in ("eval", "batch_stat"):
        m = autovc_b200.AutoEncoder(device="cuda", verbose=False, precision="bf16", batch_stat_bn=(mode == "batch_stat"))
        m.load_state_dict(sd)
        m.train(mode == "batch_stat")

        def pair():
            _, post, _ = m(x, c, c)
            return m(post, c, None)
        ms = timed(pair)
        eng = next(iter(m._engines.values()))
        eng.set_profiling(True)
        m(x, c, c); torch.cuda.synchronize(); ph1 = eng.profile()
        eng.set_profiling(False)
        out[mode] = {"ms_per_pair": round(ms, 3), "frames_per_s": round(a.B * a.T / ms * 1e3), "tflops": round(a.B * a.T * FLOP_PAIR / ms / 1e9, 1),
                     "pass1_phase_ms": {k: round(v, 4) for k, v in ph1.items()}, "launches_pass2": m.last_launch_count}
    print(json.dumps(out))


if __name__ == "__main__":
    main()
