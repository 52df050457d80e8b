"""Generate `tests/golden/*.npz` by running the UNMODIFIED reference module.  Build-container only.

Run:  python oracle/make_golden.py            (needs /root/reference; never run on the GPU box)

The reference (`/root/reference/autovc/auto_encoder/model.py:27` `AutoEncoder`) is imported with
empty stubs for the third-party modules that are not installed here (librosa, soundfile,
matplotlib, webrtcvad, noisereduce - none of them is touched by the generator forward), put in
`.eval()` under `torch.no_grad()` (the parity contract), loaded with the seeded
state dict from `autovc_b200.synth.synth_state_dict`, and run on seeded inputs.  Only seeds,
outputs and a few sub-sampled intermediate activations are stored; weights and inputs are
regenerated from the seeds by the tests.
"""
from __future__ import annotations

import os
import sys
import types

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
REF = "/root/reference"
OUT = os.path.join(ROOT, "tests", "golden")


def import_reference():
    for m in ["matplotlib", "matplotlib.pyplot", "librosa", "soundfile", "webrtcvad", "noisereduce"]:
        if m not in sys.modules:
            sys.modules[m] = types.ModuleType(m)
    sys.modules["noisereduce"].reduce_noise = lambda *a, **k: None
    os.environ.setdefault("WANDB_MODE", "disabled")
    sys.path.insert(0, REF)
    from autovc.auto_encoder.model import AutoEncoder  # noqa
    from autovc.speaker_encoder.model import SpeakerEncoder  # noqa
    return AutoEncoder, SpeakerEncoder


def mel_spec_auto_encoder_torch(wav: np.ndarray, sr: int = 22050) -> torch.Tensor:
    """Torch restatement of reference `autovc/audio/spectrogram.py:118-127` (librosa is absent):
    STFT n_fft 2048 / hop 275 / win 1100 (hann, centred, reflect pad) -> |.| -> slaney mel 80 bins
    fmin 40 -> 20*log10(max(1e-5, .)) -> clip((x+100)/100, 0, 1).  Only used to make config-1
    INPUTS (the mel front-end is before the path), so librosa bit-parity is not needed."""
    import torchaudio
    w = torch.from_numpy(wav.astype(np.float32))
    win = torch.hann_window(1100, periodic=True)
    spec = torch.stft(w, n_fft=2048, hop_length=275, win_length=1100, window=win, center=True,
                      pad_mode="reflect", return_complex=True).abs()
    fb = torchaudio.functional.melscale_fbanks(1025, 40.0, sr / 2, 80, sr, norm="slaney", mel_scale="slaney")
    mel = torch.matmul(fb.t(), spec)
    db = 20.0 * torch.log10(torch.clamp_min(mel, 1e-5))
    return torch.clamp((db + 100.0) / 100.0, 0.0, 1.0)


def normalize_volume(wav: np.ndarray, target_dbfs: float = -20.0) -> np.ndarray:
    """reference `autovc/audio/tools.py:279-282`."""
    rms = np.sqrt(np.mean(wav ** 2))
    return wav * (10 ** ((target_dbfs - 20 * np.log10(rms)) / 20))


def main():
    from autovc_b200.synth import synth_state_dict, synth_inputs
    AutoEncoder, SpeakerEncoder = import_reference()
    os.makedirs(OUT, exist_ok=True)
    torch.set_num_threads(8)

    def run_case(name, B, T, wseed=0, xseed=1234, gain=1.0, x_override=None, taps=False):
        sd = synth_state_dict(wseed, recurrent_gain=gain)
        model = AutoEncoder(device="cpu", verbose=False)
        model.load_state_dict(sd)
        model.eval()
        x, c_org, c_trg = synth_inputs(B, T, seed=xseed)
        if x_override is not None:
            x = x_override
        captured = {}
        hooks = []
        if taps:
            def mk(key, post=None):
                def hook(_m, _i, o):
                    o = o[0] if isinstance(o, tuple) else o
                    captured[key] = o.detach().clone()
                return hook
            for i in range(3):
                hooks.append(model.encoder.convolutions[i].register_forward_hook(mk(f"encoder.convbn{i}")))
                hooks.append(model.decoder.convolutions[i].register_forward_hook(mk(f"decoder.convbn{i}")))
            for i in range(5):
                hooks.append(model.postnet.convolutions[i].register_forward_hook(mk(f"postnet.convbn{i}")))
            hooks.append(model.encoder.lstm.register_forward_hook(mk("encoder.lstm")))
            hooks.append(model.decoder.lstm1.register_forward_hook(mk("decoder.lstm1")))
            hooks.append(model.decoder.lstm2.register_forward_hook(mk("decoder.lstm2")))
        with torch.no_grad():
            mel, post, codes = model(x, c_org, c_trg)
            codes_only = model(x, c_org, None)
        for h in hooks:
            h.remove()
        assert torch.equal(codes, codes_only)
        out = dict(B=B, T=T, wseed=wseed, xseed=xseed, gain=gain,
                   mel=mel.numpy(), mel_post=post.numpy(), codes=codes.numpy())
        if x_override is not None:
            out["x"] = x.numpy()
        for k, v in captured.items():
            # sub-sample: utterance 0; conv taps are (B,C,T) -> every 8th channel; lstm taps (B,T,H) -> every 8th unit
            v0 = v[0]
            out["tap." + k] = (v0[::8, :] if "conv" in k else v0[:, ::8]).numpy().copy()
        np.savez_compressed(os.path.join(OUT, name + ".npz"), **out)
        print(name, "mel|max|=%.4f post|max|=%.4f codes|max|=%.4f" % (mel.abs().max(), post.abs().max(), codes.abs().max()))
        return model, sd

    # config-1 input: real mels from two sample wavs (first 128 frames, batch 2)
    from scipy.io import wavfile
    mels = []
    for f in ["mette_183.wav", "hillary_116.wav"]:
        sr, wav = wavfile.read(os.path.join(REF, "data", "samples", f))
        assert sr == 22050 and wav.ndim == 1
        mels.append(mel_spec_auto_encoder_torch(normalize_volume(wav.astype(np.float32)))[:, :128])
    x_c1 = torch.stack(mels).contiguous()

    run_case("c1_samples_b2_t128", 2, 128, x_override=x_c1)
    run_case("synth_b2_t128", 2, 128, taps=True)
    run_case("synth_b2_t100", 2, 100)
    run_case("synth_b1_t33", 1, 33)
    run_case("synth_b1_t32", 1, 32)
    run_case("synth_b3_t400", 3, 400)
    run_case("stress_b2_t128_gain2p5", 2, 128, wseed=7, xseed=99, gain=2.5)
    model, sd = run_case("synth_b5_t64_w3", 5, 64, wseed=3, xseed=5)

    # T < freq raises in the reference; record the exception type
    x, c_org, c_trg = synth_inputs(1, 31, seed=1)
    try:
        with torch.no_grad():
            model(x, c_org, c_trg)
        raised = "none"
    except Exception as e:  # noqa
        raised = type(e).__name__
    print("T=31 raises:", raised)

    # batch_forward overlap-average: 3 chunks x 128, overlap 0.5 -> (80,256)
    x, c_org, c_trg = synth_inputs(3, 128, seed=77)
    with torch.no_grad():
        bf = model.batch_forward(x, c_org[:1], c_trg[:1], overlap=0.5)
    np.savez_compressed(os.path.join(OUT, "batch_forward_m3_n128.npz"), wseed=3, xseed=77, out=bf.numpy(),
                        t_lt_freq_raises=raised)
    print("batch_forward", tuple(bf.shape))

    # speaker encoder (d-vector): seeded default-init weights are stored via seed of torch RNG -> store outputs + weights hash
    torch.manual_seed(11)
    se = SpeakerEncoder(device="cpu", verbose=False).eval()
    rng = np.random.default_rng(5)
    se_sd = {}
    for k, v in se.state_dict().items():
        bound = 1.0 / np.sqrt(256)
        se_sd[k] = torch.from_numpy(rng.uniform(-bound, bound, size=tuple(v.shape)).astype(np.float32))
    se.load_state_dict(se_sd)
    parts = torch.from_numpy(rng.normal(size=(8, 160, 40)).astype(np.float32))
    with torch.no_grad():
        emb = se(parts)
    raw = emb.mean(dim=0)
    utt = raw / torch.linalg.norm(raw, 2)
    np.savez_compressed(os.path.join(OUT, "speaker_encoder_p8.npz"), seed=5, partial_embeds=emb.numpy(), embed=utt.numpy())
    print("speaker encoder", tuple(emb.shape))


if __name__ == "__main__":
    main()
