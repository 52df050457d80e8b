"""Where do the occasional slow end-to-end passes come from?  Repeats bench.py's e2e region and prints allocator counters."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, autovc_b200
from autovc_b200.synth import synth_inputs, synth_state_dict
from autovc_b200.voice_converter import HostPipeline
dev = torch.device("cuda", 0)
B, T, K = 256, 128, 20
m = autovc_b200.AutoEncoder(device=dev, verbose=False); m.load_state_dict(synth_state_dict(0))
x_h, co_h, ct_h = (t.pin_memory() for t in synth_inputs(B, T, seed=1))
mel_h = torch.empty(B, 80, T).pin_memory(); post_h = torch.empty(B, 80, T).pin_memory(); codes_h = torch.empty(B, 256).pin_memory()
pipe = HostPipeline(m, dev)
hin = [(x_h, co_h, ct_h)] * K; hout = [(mel_h, post_h, codes_h)] * K
pipe.run(hin[:6], hout[:6]); torch.cuda.synchronize()
for rep in range(12):
    s0 = torch.cuda.memory_stats()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter()
    a.record(); pipe.run(hin, hout); b.record()
    t1 = time.perf_counter()
    torch.cuda.synchronize()
    s1 = torch.cuda.memory_stats()
    print(f"rep {rep}: device {a.elapsed_time(b):7.2f} ms  host enqueue {1e3 * (t1 - t0):7.2f} ms  cudaMalloc {s1['num_device_alloc'] - s0['num_device_alloc']}"
          f" cudaFree {s1['num_device_free'] - s0['num_device_free']} retries {s1['num_alloc_retries'] - s0['num_alloc_retries']} reserved {s1['reserved_bytes.all.current'] >> 20} MB")
