"""Throughput of the mel front-end kernels: config-2 sized work (256 utterances x 128 frames) and one long utterance."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from autovc_b200 import audio

def timeit(fn, n=20):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(n): fn()
    e.record(); e.synchronize()
    return s.elapsed_time(e) / n

ae = audio.MelFrontend(22050, 2048, 275, 1100, 80, 40.0, power=1, log_normalize=True)
se = audio.MelFrontend(16000, 400, 160, None, 40, power=2, frames_major=True)
w = torch.rand(256, 275 * 128 - 1, device="cuda") - 0.5          # 128 frames each
ms = timeit(lambda: ae(w))
print(f"generator mel: 256 x 128 frames in {ms:.3f} ms = {256 * 128 / ms / 1e3:.2f} M frames/s; wav bytes {w.numel() * 4 / 1e6:.1f} MB")
w1 = torch.rand(22050 * 60, device="cuda") - 0.5
ms = timeit(lambda: ae(w1))
print(f"generator mel: 60 s utterance ({ae.num_frames(w1.numel())} frames) in {ms:.3f} ms")
ws = torch.rand(64 * 8, 25600, device="cuda") - 0.5               # 64 speakers x 8 partials of 1.6 s
ms = timeit(lambda: se(ws))
print(f"speaker-encoder mel: 512 x 161 frames in {ms:.3f} ms = {512 * 161 / ms / 1e3:.2f} M frames/s")
