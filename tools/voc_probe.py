"""GPU probe of the vocoder path against the oracle: conditioning, a teacher-forced run, a free run; then timings."""
import os, sys, time
import numpy as np
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from autovc_b200.vocoder import WaveRNN, fold_params
from autovc_b200.synth import synth_wavernn_state_dict, synth_uniforms
from oracle import wavernn_oracle as wo

sd = synth_wavernn_state_dict(11)
voc = WaveRNN(device="cuda")
voc.load_state_dict(sd)
T, target, overlap = 24, 1500, 100
mel = np.random.Generator(np.random.PCG64(3)).uniform(0, 1, size=(1, 80, T)).astype(np.float32)
m_ref, aux_ref = wo.upsample(sd, np.pad(mel, ((0, 0), (0, 0), (2, 2))))
m_up, aux = voc.condition(torch.from_numpy(mel).cuda())
torch.cuda.synchronize()
print("cond m err", np.abs(m_up.cpu().numpy() - m_ref).max(), "aux err", np.abs(aux.cpu().numpy() - aux_ref[:, ::275]).max())
mels_f, aux_f, wave_len = wo.condition(sd, mel, True, target, overlap)
B, L = mels_f.shape[:2]
n = int(os.environ.get("N", 300))
U = synth_uniforms(L, B, 21)[:n]
free_ref, lg_ref = wo.generate_rows(sd, mels_f[:, :n], aux_f[:, :n], U, return_logits=True)
starts = [i * (target + overlap) for i in range(B)]
t0 = time.time()
s_forced, lg = voc.sample_rows(m_up, aux, [0] * B, starts, n, uniforms=U, forced_x=free_ref.T.copy(), return_logits=True)
torch.cuda.synchronize()
print("forced run", time.time() - t0, "s")
lg = lg.cpu().numpy(); s_forced = s_forced.cpu().numpy()
e = np.abs(lg - lg_ref)
print("forced logits err max", e.max(), "mean", e.mean(), "per-step max (first 8)", e.max(axis=(1, 2))[:8])
print("forced sample err max", np.abs(s_forced - free_ref).max(), "mean", np.abs(s_forced - free_ref).mean())
s_free = voc.sample_rows(m_up, aux, [0] * B, starts, n, uniforms=U).cpu().numpy()
d = np.abs(s_free - free_ref).max(axis=0)
print("free sample err: first 10", d[:10], "max over", n, d.max(), "steps with err>0.05:", int((d > 0.05).sum()))
for rows, steps in ((5, 2000), (128, 2000), (256, 2000), (512, 1000)):
    mm = torch.rand(1, 80, 64, device="cuda")
    mu, ax = voc.condition(mm)
    ru = [0] * rows; rs = [(i * 97) % 4000 for i in range(rows)]
    voc.sample_rows(mu, ax, ru, rs, 50)
    torch.cuda.synchronize()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record(); out = voc.sample_rows(mu, ax, ru, rs, steps); ev1.record(); torch.cuda.synchronize()
    ms = ev0.elapsed_time(ev1)
    print(f"rows {rows} steps {steps}: {ms:.1f} ms, {ms / steps * 1e3:.2f} us/step, {rows * steps / ms / 1e3:.2f} M samples/s, finite {bool(torch.isfinite(out).all())} std {float(out.std()):.3f}")
