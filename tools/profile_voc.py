"""One short vocoder loop for ncu: 512 rows x STEPS steps (default 200) after one warm-up call."""
import os, sys
import numpy as np
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from autovc_b200.vocoder import WaveRNN
from autovc_b200.synth import synth_wavernn_state_dict

voc = WaveRNN(device="cuda")
voc.load_state_dict(synth_wavernn_state_dict(11))
m_up, aux = voc.condition(torch.rand(4, 80, 64, device="cuda"))
rows, steps = int(os.environ.get("ROWS", 512)), int(os.environ.get("STEPS", 200))
ru = (np.arange(rows) % 4).astype(np.int32)
rs = ((np.arange(rows) * 97) % (60 * 275)).astype(np.int32)
for _ in range(2):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); s = voc.sample_rows(m_up, aux, ru, rs, steps); e1.record(); torch.cuda.synchronize()
print(f"rows {rows} steps {steps}: {e0.elapsed_time(e1):.2f} ms, launches {voc.last_launch_count}, finite {bool(torch.isfinite(s).all())}")
