"""Tiny end-to-end case for compute-sanitizer (one tool per gpurun call): a bf16 generator forward (B=3, T=64: tc_gemm,
enc_lstm, ll_lstm H=512/1024), the speaker encoder (3 partials x 40 frames: ll_lstm H=256) and the vocoder loop (3 rows x 24
steps).  Persistent kernels are launched without the cooperative attribute (AVC_LSTM_NOCOOP / AVC_VOC_NOCOOP): the
sanitizer replaces the launch path; the grids are co-resident either way."""
import os, sys
os.environ.setdefault("AVC_LSTM_NOCOOP", "1")
os.environ.setdefault("AVC_VOC_NOCOOP", "1")
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import autovc_b200
from autovc_b200.synth import synth_inputs, synth_speaker_encoder, synth_state_dict, synth_uniforms, synth_wavernn_state_dict

what = sys.argv[1:] or ["gen", "se", "voc"]
if "gen" in what:
    m = autovc_b200.AutoEncoder(device="cuda", verbose=False)
    m.load_state_dict(synth_state_dict(0))
    x, c1, c2 = (t.cuda() for t in synth_inputs(3, 64, seed=4))
    mel, post, codes = m(x, c1, c2)
    torch.cuda.synchronize()
    print("generator:", tuple(post.shape), float(post.abs().max()), "launches", m.last_launch_count)
if "se" in what:
    sd, parts = synth_speaker_encoder(5, n_partials=3, n_frames=40)
    se = autovc_b200.SpeakerEncoder(device="cuda", verbose=False)
    se.load_state_dict(sd)
    e = se(parts.cuda())
    torch.cuda.synchronize()
    print("speaker encoder:", tuple(e.shape), float(e.norm(dim=1).mean()))
if "voc" in what:
    vsd = synth_wavernn_state_dict(11)
    voc = autovc_b200.WaveRNN(device="cuda")
    voc.load_state_dict(vsd)
    mel = torch.rand(1, 80, 12, generator=torch.Generator().manual_seed(3)).cuda()
    m_up, aux = voc.condition(mel)
    n = 24
    U = synth_uniforms(n, 3, 1)
    got = voc.sample_rows(m_up, aux, [0, 0, 0], [0, 300, 900], n, uniforms=U)
    torch.cuda.synchronize()
    print("vocoder:", tuple(got.shape), float(got.abs().max()))
print("done")
