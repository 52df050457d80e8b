"""Generate `tests/golden/wavernn_*.npz` by running the UNMODIFIED reference `WaveRNN` (`autovc/wavernn/model.py:95`).

Run:  python oracle/make_golden_wavernn.py     (build container only: needs /root/reference)

The module is imported as in make_golden.py (empty stubs for absent third-party modules; `numpy.cumproduct`, removed in
numpy 2, aliased to `cumprod` - an environment shim, the reference file is untouched), loaded with
`autovc_b200.synth.synth_wavernn_state_dict`, and `generate` runs with `torch.Tensor.uniform_` patched to read the
explicit stream `synth_uniforms` in call order (10 mixture uniforms, then 1 logistic uniform per step and row), which
is how the oracle and the CUDA path consume it.  Stored: seeds, the waveform, the up-sampled conditioning sub-sampled,
and the raw per-fold samples.
"""
from __future__ import annotations

import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))
OUT = os.path.join(ROOT, "tests", "golden")


def main():
    np.cumproduct = np.cumprod
    import make_golden
    make_golden.import_reference()
    from autovc.wavernn.model import WaveRNN
    import autovc.wavernn.model as wm
    from autovc_b200.synth import synth_wavernn_state_dict, synth_uniforms

    wm.WaveRNN.gen_display = lambda *a, **k: None
    torch.set_num_threads(8)

    def run(name, T, target, overlap, batched, wseed=11, mseed=3, useed=21):
        sd = synth_wavernn_state_dict(wseed)
        model = WaveRNN(device="cpu")
        model.load_state_dict(sd)
        rng = np.random.Generator(np.random.PCG64(mseed))
        mel = rng.uniform(0, 1, size=(1, 80, T)).astype(np.float32)
        # rows / steps exactly as generate derives them
        total = T * 275
        if batched:
            nf = (total - overlap) // (target + overlap)
            if total - (nf * (overlap + target) + overlap) != 0:
                nf += 1
            B, L = nf, target + 2 * overlap
        else:
            B, L = 1, total
        U = synth_uniforms(L, B, useed)
        state = {"t": 0, "phase": 0}
        taps = {}

        def fake_uniform(self, a=0.0, b=1.0, **kw):
            if kw or (a, b) != (1e-5, 1.0 - 1e-5):               # weight initialisers of the GRUCell copies (model.py:336)
                return orig(self, a, b, **kw)
            t = state["t"]
            if state["phase"] == 0:
                assert tuple(self.shape) == (1, B, 10), self.shape
                self.copy_(torch.from_numpy(U[t, :, :10]).view(1, B, 10))
                state["phase"] = 1
            else:
                assert tuple(self.shape) == (1, B), self.shape
                self.copy_(torch.from_numpy(U[t, :, 10]).view(1, B))
                state["phase"] = 0
                state["t"] = t + 1
            return self

        up = {}
        hook = model.upsample.register_forward_hook(lambda _m, _i, o: up.update(m=o[0].detach().clone(), aux=o[1].detach().clone()))
        stack = torch.stack

        def tap_stack(xs, *a, **k):
            r = stack(xs, *a, **k)
            taps["samples"] = r.detach().clone()
            return r

        orig = torch.Tensor.uniform_
        torch.Tensor.uniform_ = fake_uniform
        wm.torch.stack = tap_stack
        try:
            wav = model.generate(torch.from_numpy(mel), batched=batched, target=target, overlap=overlap, mu_law=False)
        finally:
            torch.Tensor.uniform_ = orig
            wm.torch.stack = stack
            hook.remove()
        assert state["t"] == L, (state, L)
        samples = taps["samples"].numpy().T.copy()                 # (B, L)
        np.savez_compressed(os.path.join(OUT, name + ".npz"), T=T, target=target, overlap=overlap, batched=batched, wseed=wseed,
                            mseed=mseed, useed=useed, wav=wav, samples=samples,
                            up_m=up["m"].numpy()[0, ::37, ::7].copy(), up_aux=up["aux"].numpy()[0, ::275, ::5].copy())
        print(name, "rows", B, "steps", L, "wav", wav.shape, "|wav|max %.3f" % np.abs(wav).max(), "std %.3f" % samples.std())

    run("wavernn_t24_folded", 24, 1500, 100, True)
    run("wavernn_t22_single", 22, 0, 0, False)
    run("wavernn_t22_exactfold", 22, 2950, 50, True)        # 6050 = 2 * 3000 + 50: no padded fold


if __name__ == "__main__":
    main()
