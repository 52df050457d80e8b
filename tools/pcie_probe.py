"""Under torchrun: per-rank H2D / D2H times of the config-2 step payload (11 MB in, 21 MB out), alone vs all ranks at once,
and the HostPipeline e2e rate for several step counts.  Diagnoses the N=8 end-to-end efficiency."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.distributed as dist
import autovc_b200
from autovc_b200.synth import synth_inputs, synth_state_dict
from autovc_b200.voice_converter import HostPipeline

rank, world, lr = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
torch.cuda.set_device(lr); dev = torch.device("cuda", lr)
def barrier():
    torch.cuda.synchronize()
    if world > 1: dist.barrier()
    torch.cuda.synchronize()
hin = torch.empty(11 << 20, dtype=torch.uint8).pin_memory(); hout = torch.empty(21 << 20, dtype=torch.uint8).pin_memory()
din = torch.empty(11 << 20, dtype=torch.uint8, device=dev); dout = torch.empty(21 << 20, dtype=torch.uint8, device=dev)
def t_copy(fn, reps=20):
    fn(); torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(reps): fn()
    e.record(); e.synchronize()
    return s.elapsed_time(e) / reps
res = {}
for mode in ("alone", "together"):
    for name, fn in (("h2d", lambda: din.copy_(hin, non_blocking=True)), ("d2h", lambda: hout.copy_(dout, non_blocking=True))):
        if mode == "alone":
            for r in range(world):
                barrier()
                if r == rank: res[f"{name}_alone"] = t_copy(fn)
        else:
            barrier(); res[f"{name}_together"] = t_copy(fn)
    if mode == "together":   # both directions at once, all ranks
        barrier()
        s2 = torch.cuda.Stream()
        def both():
            din.copy_(hin, non_blocking=True)
            with torch.cuda.stream(s2): hout.copy_(dout, non_blocking=True)
        t = t_copy(both); torch.cuda.synchronize(); res["both_together"] = t
model = autovc_b200.AutoEncoder(device=dev, verbose=False); model.load_state_dict(synth_state_dict(0))
x_h, co_h, ct_h = (t.pin_memory() for t in synth_inputs(256, 128, seed=1 + rank))
outs = (torch.empty(256, 80, 128).pin_memory(), torch.empty(256, 80, 128).pin_memory(), torch.empty(256, 256).pin_memory())
pipe = HostPipeline(model, dev)
pipe.run([(x_h, co_h, ct_h)] * 4, [outs] * 4); barrier()
xd, cd, td = x_h.to(dev), co_h.to(dev), ct_h.to(dev)
for _ in range(3): model(xd, cd, td)
barrier(); s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
s.record()
for _ in range(20): model(xd, cd, td)
e.record(); barrier(); res["dev_ms_step"] = s.elapsed_time(e) / 20
for K in (20, 60):
    barrier(); s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter(); s.record(); pipe.run([(x_h, co_h, ct_h)] * K, [outs] * K); e.record(); t1 = time.perf_counter(); barrier()
    res[f"e2e_ms_step_K{K}"] = s.elapsed_time(e) / K; res[f"enqueue_ms_step_K{K}"] = (t1 - t0) * 1e3 / K
print(f"rank {rank}: " + "  ".join(f"{k} {v:.3f}" for k, v in res.items()), flush=True)
if world > 1: dist.barrier(); dist.destroy_process_group()
