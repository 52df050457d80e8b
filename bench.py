#!/usr/bin/env python
"""Benchmark of the generator forward pass (BASELINE.json metric: mel-frames/s; p50 latency per 128-frame utt).

    python bench.py [--gpus N] [--steps K] [--warmup W]            # this repo's CUDA path
    python bench.py --impl reference [...]                         # the reference algorithm on the host CPU cores

One "step" = one `AutoEncoder.forward(x, c_org, c_trg)` over one synthetic batch of config 2
(B=256 utterances x T=128 frames per GPU, default dims 32/256/512/32, seeded random-init weights, bf16 mode).
N>1 (launched by torchrun, one rank per GPU): utterances are independent, so every rank runs its own batch with
replicated weights and there is no collective on the data path ("weak" scaling); the timed region is bracketed by
a barrier + synchronize and the MAX over ranks is reported.
Prints ONE JSON line (rank 0).
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

FLOP_PER_FRAME = 56_770_560            # 2 x 28,385,280 MAC, every weight used once per frame
MAC_PER_FRAME = {                      # per phase
    "conv": 3_481_600 + 3_932_160 + 4_341_760,
    "proj": 147_456 + 655_360 + 6_291_456 + 81_920,
    "enc_lstm": 16_384,
    "lstm1": 1_048_576,
    "lstm2": 8_388_608,
}
REALTIME_FRAMES_PER_S = 22050 / 275    # reference utils/hparams.py:7,10


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            d = json.load(f)
        return d, "measured"
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0}, "fallback"


class ClockSampler:
    """SM clock and throttle reasons of one GPU, sampled every 5 ms through NVML while the timed region runs (the region is
    tens of milliseconds: an `nvidia-smi -lms` child process does not even start in that time on an 8-GPU box); falls back to
    the nvidia-smi loop when the NVML binding is missing."""

    FIELDS = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
              "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
              "clocks_event_reasons.sw_power_cap")
    REASONS = ((0x8, "hw_slowdown"), (0x40, "hw_thermal_slowdown"), (0x20, "sw_thermal_slowdown"), (0x4, "sw_power_cap"))

    def __init__(self, index: int):
        self.index = index
        self.rows = []
        self.samples = []          # (sm_mhz, reasons bitmask) from NVML
        self.proc = None
        self.thread = None
        self.nvml = None
        self.handle = None
        self.max_mhz = None
        self._stop = threading.Event()
        try:
            import pynvml
            pynvml.nvmlInit()
            try:
                uuid = str(torch.cuda.get_device_properties(index).uuid)
                self.handle = pynvml.nvmlDeviceGetHandleByUUID(("GPU-" + uuid) if not uuid.startswith("GPU-") else uuid)
            except Exception:
                self.handle = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = float(pynvml.nvmlDeviceGetMaxClockInfo(self.handle, pynvml.NVML_CLOCK_SM))
            self.nvml = pynvml
        except Exception:
            self.nvml = None

    def start(self):
        if self.nvml is not None:
            nv, h = self.nvml, self.handle

            def poll():
                while not self._stop.is_set():
                    try:
                        self.samples.append((float(nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM)),
                                             int(nv.nvmlDeviceGetCurrentClocksThrottleReasons(h))))
                    except Exception:
                        pass
                    time.sleep(0.005)

            self.thread = threading.Thread(target=poll, daemon=True)
            self.thread.start()
            return
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.FIELDS}", "--format=csv,noheader,nounits", "-lms", "100"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except Exception:
            self.proc = None
            return

        def pump():
            for line in self.proc.stdout:
                self.rows.append(line.strip())

        self.thread = threading.Thread(target=pump, daemon=True)
        self.thread.start()

    def stop(self):
        if self.nvml is not None:
            self._stop.set()
            if self.thread is not None:
                self.thread.join(timeout=2)
            sm = [s for s, _ in self.samples]
            mask = 0
            for _, m in self.samples:
                mask |= m
            return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": self.max_mhz,
                    "reasons": sorted(n for bit, n in self.REASONS if mask & bit), "samples": len(sm), "source": "nvml, 5 ms period"}
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        for r in self.rows:
            parts = [p.strip() for p in r.split(",")]
            if len(parts) < 7:
                continue
            try:
                sm.append(float(parts[0]))
                mx.append(float(parts[1]))
            except ValueError:
                continue
            for name, val in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), parts[3:7]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm), "source": "nvidia-smi -lms 100"}


def time_oracle(batch: int, frames: int, steps: int, warmup: int):
    """The reference algorithm (oracle port, torch CPU fp32) on the host cores: frames/s."""
    from oracle import autovc_oracle as O
    from autovc_b200.synth import synth_inputs, synth_state_dict
    torch.set_num_threads(os.cpu_count() or 1)
    sd = synth_state_dict(0)
    x, c_org, c_trg = synth_inputs(batch, frames, seed=1234)
    times = []
    for i in range(warmup + steps):
        t0 = time.perf_counter()
        O.generator_forward(sd, x, c_org, c_trg)
        dt = time.perf_counter() - t0
        if i >= warmup:
            times.append(dt)
    total = sum(times)
    return batch * frames * len(times) / total, total / len(times) * 1e3, torch.get_num_threads()


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    sample_b = args.cpu_sample_batch
    fps, ms, cores = time_oracle(sample_b, args.frames, args.steps, args.warmup)
    sample = f"{sample_b} of the {args.batch} utterances x {args.frames} frames per step (oracle port of the reference algorithm, torch CPU fp32, all host threads)"
    line = {
        "impl": "reference", "metric": "generator_fwd_mel_frames_per_s", "value": fps, "unit": "frames/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"config2: B={args.batch} x T={args.frames} x 80 mel, dims 32/256/512/32, seeded random-init", "sample": sample},
        "cpu_baseline": {"value": fps, "unit": "frames/s", "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": fps, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--batch", type=int, default=256, help="utterances per GPU per step (config 2: 256)")
    ap.add_argument("--frames", type=int, default=128)
    ap.add_argument("--precision", default="bf16", choices=["bf16", "fp32"])
    ap.add_argument("--cpu-sample-batch", type=int, default=16)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup

    if args.impl == "reference":
        run_reference_arm(args)
        return

    import autovc_b200
    from autovc_b200.synth import synth_inputs, synth_state_dict

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    dist = None
    if world > 1:
        import torch.distributed as dist
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    B, T, K, W = args.batch, args.frames, args.steps, args.warmup

    model = autovc_b200.AutoEncoder(device=dev, verbose=False, precision=args.precision)
    model.load_state_dict(synth_state_dict(0))
    x_h, co_h, ct_h = synth_inputs(B, T, seed=1234 + rank)
    x_h, co_h, ct_h = x_h.pin_memory(), co_h.pin_memory(), ct_h.pin_memory()
    x, c_org, c_trg = x_h.to(dev), co_h.to(dev), ct_h.to(dev)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)   # > 126 MB L2

    def barrier():
        torch.cuda.synchronize(dev)
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize(dev)

    # ---- warm-up (also builds the engine, packs weights, creates tensor maps) ----
    for _ in range(W):
        model(x, c_org, c_trg)
    eng = model._engine(dev)
    launches_per_step = eng.last_launch_count

    # ---- device-timed region: inputs resident in HBM ----
    sampler = ClockSampler(local_rank)
    starts = [torch.cuda.Event(enable_timing=True) for _ in range(K)]
    ends = [torch.cuda.Event(enable_timing=True) for _ in range(K)]
    barrier()
    sampler.start()
    wall0 = time.perf_counter()
    for i in range(K):
        flush.zero_()                      # evict L2 between timed iterations (not timed)
        starts[i].record()
        model(x, c_org, c_trg)
        ends[i].record()
    barrier()
    wall = time.perf_counter() - wall0
    clocks = sampler.stop()
    dev_ms = sum(s.elapsed_time(e) for s, e in zip(starts, ends))

    # ---- end-to-end through the public API with HOST buffers (pinned), copies inside the timed region ----
    mel_h = torch.empty(B, 80, T, dtype=torch.float32).pin_memory()
    post_h = torch.empty(B, 80, T, dtype=torch.float32).pin_memory()
    n_codes = 32 * (T // 32 + (T + 31) // 32)
    codes_h = torch.empty(B, n_codes, dtype=torch.float32).pin_memory()

    # HostPipeline (autovc_b200/voice_converter.py): H2D of step i+1 and D2H of step i-1 ride a copy stream while
    # step i computes; every step still copies its own inputs in and all three outputs out inside the timed region
    from autovc_b200.voice_converter import HostPipeline
    pipe = HostPipeline(model, dev)
    host_in = [(x_h, co_h, ct_h)] * K
    host_out = [(mel_h, post_h, codes_h)] * K
    pipe.run(host_in[:max(2, min(K, 6))], host_out[:max(2, min(K, 6))])   # untimed: fills the allocator's pools for both streams
    barrier()
    e_start, e_end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    flush.zero_()
    barrier()
    e_start.record()
    pipe.run(host_in, host_out)
    e_end.record()
    barrier()
    e2e_ms = e_start.elapsed_time(e_end)
    h2d = x_h.numel() * 4 + co_h.numel() * 4 + ct_h.numel() * 4
    d2h = mel_h.numel() * 4 + post_h.numel() * 4 + codes_h.numel() * 4

    # ---- per-phase device time (CUDA events inside the library, same stream) for the roofline ----
    eng.set_profiling(True)
    phase_ms = {}
    PROF_STEPS = min(K, 5)
    for _ in range(PROF_STEPS):
        flush.zero_()
        model(x, c_org, c_trg)
        for k, v in eng.profile().items():
            phase_ms[k] = phase_ms.get(k, 0.0) + v / PROF_STEPS
    eng.set_profiling(False)

    # ---- p50 latency of a single 128-frame utterance (B=1) ----
    x1, co1, ct1 = x[:1].contiguous(), c_org[:1].contiguous(), c_trg[:1].contiguous()
    for _ in range(3):
        model(x1, co1, ct1)
    lat = []
    for _ in range(30):
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        model(x1, co1, ct1)
        e.record()
        e.synchronize()
        lat.append(s.elapsed_time(e))
    lat_p50 = statistics.median(lat)

    # ---- reduce over ranks (max time) ----
    t = torch.tensor([dev_ms, e2e_ms], dtype=torch.float64, device=dev)
    if dist is not None:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    dev_ms, e2e_ms = float(t[0]), float(t[1])

    if rank == 0:
        peaks, peak_src = load_peaks()
        frames_total = world * B * T * K
        value = frames_total / (dev_ms * 1e-3)
        e2e_value = frames_total / (e2e_ms * 1e-3)
        dom = max((k for k in phase_ms if k in MAC_PER_FRAME), key=lambda k: phase_ms[k])
        dom_flops = 2.0 * MAC_PER_FRAME[dom] * B * T
        achieved = dom_flops / (phase_ms[dom] * 1e-3) / 1e12 if phase_ms[dom] > 0 else 0.0
        peak = float(peaks.get("bf16_tflops_sustained", peaks.get("bf16_tflops")))
        # DRAM bytes of the dominant phase's launches from the committed ncu --set full capture (same config), or null
        traffic = None
        try:
            tj = json.load(open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "profiles", "r1_ncu_traffic.json")))
            if (B, T) == (256, 128) and args.precision == "bf16":
                traffic = tj["bytes_per_step"].get(dom)
        except Exception:
            traffic = None
        line = {
            "metric": "generator_fwd_mel_frames_per_s", "value": value, "unit": "frames/s", "n_gpus": world,
            "steps": K, "warmup": W, "ms_per_step": dev_ms / K, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "bf16" if args.precision == "bf16" else "f32", "data": "synthetic",
            "config": {"workload": f"config2: B={B} x T={T} x 80 mel per GPU, dims 32/256/512/32, seeded random-init weights",
                       "l2": "256 MB flush written between timed iterations (excluded from the timing)",
                       "parallelism": f"replicas x{world}, no data-path collective",
                       "e2e": "HostPipeline: pinned host buffers, per-step H2D + D2H on a copy stream overlapped with the previous/next step's compute"},
            "realtime_x": value / REALTIME_FRAMES_PER_S,
            "whole_path_tensor_frac": value / world * FLOP_PER_FRAME / (peak * 1e12),
            "latency_p50_ms_b1_t128": lat_p50,
            "e2e": {"value": e2e_value, "unit": "frames/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h},
            "gpu_launches": int(launches_per_step * K),
            "clocks": clocks,
            "phase_ms": {k: round(v, 4) for k, v in phase_ms.items()},
            "roofline": {"bound": "tensor", "kernel": dom, "achieved": achieved, "peak": peak, "unit": "TFLOP/s",
                         "frac": achieved / peak, "traffic": traffic, "peak_source": f"{peak_src} bf16_tflops_sustained",
                         "algorithmic_flop_per_launch_group": dom_flops},
            "wall_s_timed_region": wall,
        }
        if not args.no_cpu_baseline and world == 1:     # the CPU leg is an N=1 measurement (rank 0 would share the host with the other ranks)
            fps, ms, cores = time_oracle(args.cpu_sample_batch, T, 3, 1)
            line["cpu_baseline"] = {"value": fps, "unit": "frames/s", "cores": cores, "kind": "port",
                                    "sample": f"{args.cpu_sample_batch} of the {B} utterances x {T} frames, 3 timed passes of the oracle port (torch CPU fp32)"}
        print(json.dumps(line))
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
