"""Timing probe of config 4's pieces (speaker encoder batch, convert_multiple) - repeated calls, wall and events."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import autovc_b200
from autovc_b200.synth import synth_state_dict, synth_speaker_encoder
from autovc_b200.voice_converter import Utterance, VoiceConverter

dev = torch.device("cuda", 0)
model = autovc_b200.AutoEncoder(device=dev, verbose=False); model.load_state_dict(synth_state_dict(0))
se_sd, _ = synth_speaker_encoder(5)
se = autovc_b200.SpeakerEncoder(device=dev, verbose=False); se.load_state_dict(se_sd)
N4, S4, T4 = 4096, 64, 128
g = torch.Generator().manual_seed(4242)
partials = [torch.randn(8, 160, 40, generator=g) for _ in range(S4)]
slab = torch.rand(N4, 80, T4).pin_memory()
dummy = torch.zeros(80, 1)
srcs = [Utterance(slab[i], f"spk{i * S4 // N4}", partials[i * S4 // N4]) for i in range(N4)]
trgs = [Utterance(dummy, f"spk{(i * S4 // N4 + 1) % S4}", partials[(i * S4 // N4 + 1) % S4]) for i in range(N4)]
vc = VoiceConverter(model, embed_speaker=se.as_embed_fn(), device=dev, max_batch=256)
keys = [u.speaker for u in srcs] + [u.speaker for u in trgs]
datas = [u.speaker_data for u in srcs] + [u.speaker_data for u in trgs]
for rep in range(4):
    vc.speakers.clear()
    torch.cuda.synchronize(); t0 = time.perf_counter()
    vc.learn_speakers(keys, datas, embed_many=se.embed_speakers)
    torch.cuda.synchronize(); t1 = time.perf_counter()
    print(f"learn_speakers rep {rep}: {1e3 * (t1 - t0):.2f} ms")
flat = torch.cat(partials).to(dev)
for rep in range(3):
    torch.cuda.synchronize(); t0 = time.perf_counter()
    se(flat); torch.cuda.synchronize(); t1 = time.perf_counter()
    print(f"se forward 512x160 rep {rep}: {1e3 * (t1 - t0):.2f} ms")
for rep in range(4):
    torch.cuda.synchronize(); t0 = time.perf_counter()
    res = vc.convert_multiple(srcs, trgs, match_method="align", output="host")
    torch.cuda.synchronize(); t1 = time.perf_counter()
    print(f"convert_multiple rep {rep}: {1e3 * (t1 - t0):.2f} ms -> {N4 * T4 / (t1 - t0) / 1e6:.2f} M frames/s")
    del res
import cProfile, pstats
pr = cProfile.Profile(); pr.enable()
res = vc.convert_multiple(srcs, trgs, match_method="align", output="host")
pr.disable()
pstats.Stats(pr).sort_stats("cumulative").print_stats(18)
