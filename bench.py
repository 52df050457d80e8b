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


def _reference_module(device: str = "cpu"):
    """The UNMODIFIED reference `AutoEncoder` (reference autovc/auto_encoder/model.py:27), `.eval()`, seeded weights, on
    `device`; None when neither /root/reference nor the staged copy oracle/_ref is reachable."""
    try:
        from oracle.ref_loader import import_reference
        AutoEncoder, _SE, root = import_reference(prefer_staged=True)
    except Exception as exc:      # noqa: BLE001 - reported in the JSON line
        return None, f"{type(exc).__name__}: {exc}"
    from autovc_b200.synth import synth_state_dict
    model = AutoEncoder(device=device, verbose=False)
    model.load_state_dict(synth_state_dict(0))
    model.eval()
    return model.to(device), root


def time_cpu_reference(batch: int, frames: int, steps: int, warmup: int, budget_s: float = 0.0):
    """frames/s of the reference forward on the host cores: the unmodified reference module (kind "reference") when it is
    reachable, else the oracle port of its algorithm (kind "port").  `budget_s` > 0 shrinks the per-step sample (fewer
    utterances of the same batch) when a first probe step says the full batch would not fit the budget."""
    from autovc_b200.synth import synth_inputs, synth_state_dict
    torch.set_num_threads(os.cpu_count() or 1)
    model, where = _reference_module("cpu")
    x, c_org, c_trg = synth_inputs(batch, frames, seed=1234)
    if model is not None:
        kind = "reference"

        def step(n):
            with torch.no_grad():
                model(x[:n], c_org[:n], c_trg[:n])
    else:
        from oracle import autovc_oracle as O
        sd = synth_state_dict(0)
        kind = "port"

        def step(n):
            O.generator_forward(sd, x[:n], c_org[:n], c_trg[:n])
    n = batch
    t0 = time.perf_counter()
    step(min(batch, 16))                                   # probe (also the first warm-up: thread pools, allocator)
    probe = (time.perf_counter() - t0) * batch / min(batch, 16)
    if budget_s > 0 and probe * (steps + warmup) > budget_s:
        n = max(16, int(batch * budget_s / (probe * (steps + warmup))) // 16 * 16)
    times = []
    for i in range(warmup + steps):
        t0 = time.perf_counter()
        step(n)
        dt = time.perf_counter() - t0
        if i >= warmup:
            times.append(dt)
    total = sum(times)
    sample = (f"{n} of the {batch} utterances x {frames} frames per step, {len(times)} timed steps; "
              + ("the unmodified reference AutoEncoder.eval() under no_grad (torch CPU fp32, all host threads), imported from " + str(where)
                 if kind == "reference" else f"oracle port of the reference algorithm (torch CPU fp32; reference module unavailable: {where})"))
    return {"value": n * frames * len(times) / total, "unit": "frames/s", "cores": torch.get_num_threads(), "kind": kind,
            "sample": sample, "ms_per_step": total / len(times) * 1e3, "utterances_per_step": n}


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    # the whole K + W run is bounded to ~6 minutes of host time; on the B200 hosts the full batch fits (about 3 s per step)
    r = time_cpu_reference(args.batch, args.frames, args.steps, args.warmup, budget_s=360.0)
    fps = r["value"]
    line = {
        "impl": "reference", "metric": "generator_fwd_mel_frames_per_s", "value": fps, "unit": "frames/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": r["ms_per_step"], "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"config2: B={args.batch} x T={args.frames} x 80 mel per GPU, dims 32/256/512/32, seeded random-init weights",
                   "sample": r["sample"], "same_config": r["utterances_per_step"] == args.batch},
        "cpu_baseline": {k: r[k] for k in ("value", "unit", "cores", "kind", "sample")},
        "e2e": {"value": fps, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))


def gpu_eager_baseline(dev, B, T, x, c_org, c_trg):
    """The same-box bar: the unmodified reference module `.eval().cuda()` through
    PyTorch eager + cuDNN/cuBLAS on this B200 - fp32 with TF32 off, and under autocast(bfloat16) - at config 2 and at
    B = 1.  Returns milliseconds per forward (CUDA events, median)."""
    model, where = _reference_module("cpu")
    if model is None:
        return {"unavailable": where}
    model = model.to(dev)
    model.device = dev
    tf32 = (torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32)
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False

    def timed(fn, warm, reps):
        for _ in range(warm):
            fn()
        ts = []
        for _ in range(reps):
            s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            s.record(); fn(); e.record(); e.synchronize()
            ts.append(s.elapsed_time(e))
        return statistics.median(ts)

    out = {"module": f"unmodified reference AutoEncoder.eval().cuda() from {where}, torch {torch.__version__} eager (cuDNN LSTM/conv)",
           "config": f"B={B} x T={T}"}
    try:
        with torch.no_grad():
            x1, co1, ct1 = x[:1].contiguous(), c_org[:1].contiguous(), c_trg[:1].contiguous()
            out["fp32_ms"] = timed(lambda: model(x, c_org, c_trg), 3, 10)
            out["b1_fp32_ms"] = timed(lambda: model(x1, co1, ct1), 5, 30)
            with torch.autocast("cuda", dtype=torch.bfloat16):
                out["bf16_ms"] = timed(lambda: model(x, c_org, c_trg), 3, 10)
                out["b1_bf16_ms"] = timed(lambda: model(x1, co1, ct1), 5, 30)
            out["fp32_frames_per_s"] = B * T / (out["fp32_ms"] * 1e-3)
            out["bf16_frames_per_s"] = B * T / (out["bf16_ms"] * 1e-3)
            out["b1_ms"] = min(out["b1_fp32_ms"], out["b1_bf16_ms"])
    except Exception as exc:      # noqa: BLE001
        out["error"] = f"{type(exc).__name__}: {exc}"
    finally:
        torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32 = tf32
        del model
        torch.cuda.empty_cache()
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--batch", type=int, default=256, help="utterances per GPU per step (config 2: 256)")
    ap.add_argument("--frames", type=int, default=128)
    ap.add_argument("--precision", default="bf16", choices=["bf16", "fp32"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-configs", action="store_true", help="skip the extra configs (c3/c4/c5/b1) and the cuDNN-eager baseline")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup

    if args.impl == "reference":
        run_reference_arm(args)
        return

    import autovc_b200
    from autovc_b200.synth import synth_inputs, synth_state_dict
    from autovc_b200.voice_converter import HostPipeline, Utterance, VoiceConverter

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

    sd = synth_state_dict(0)
    model = autovc_b200.AutoEncoder(device=dev, verbose=False, precision=args.precision)
    model.load_state_dict(sd)
    x_h, co_h, ct_h = synth_inputs(B, T, seed=1234 + rank)
    x_h, co_h, ct_h = x_h.pin_memory(), co_h.pin_memory(), ct_h.pin_memory()
    x, c_org, c_trg = x_h.to(dev), co_h.to(dev), ct_h.to(dev)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)   # > 126 MB L2

    def barrier():
        torch.cuda.synchronize(dev)
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize(dev)

    def max_over_ranks(*vals):
        t = torch.tensor(vals, dtype=torch.float64, device=dev)
        if dist is not None:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return [float(v) for v in t]

    def timed_steps(fn, warm, steps, flush_l2=True):
        """Sum of the device time (CUDA events on the launching stream) of `steps` calls of fn, bracketed by barriers."""
        for _ in range(warm):
            fn()
        ss = [torch.cuda.Event(enable_timing=True) for _ in range(steps)]
        es = [torch.cuda.Event(enable_timing=True) for _ in range(steps)]
        barrier()
        for i in range(steps):
            if flush_l2:
                flush.zero_()                  # evict L2 between timed iterations (not timed)
            ss[i].record()
            fn()
            es[i].record()
        barrier()
        return sum(s.elapsed_time(e) for s, e in zip(ss, es))

    # ---- warm-up (also builds the engine, packs weights, creates tensor maps) ----
    for _ in range(W):
        model(x, c_org, c_trg)
    eng = model._engine(dev)
    launches_per_step = eng.last_launch_count

    # ---- device-timed region: inputs resident in HBM ----
    sampler = ClockSampler(local_rank)
    barrier()
    sampler.start()
    wall0 = time.perf_counter()
    dev_ms = timed_steps(lambda: model(x, c_org, c_trg), 0, K)
    wall = time.perf_counter() - wall0
    clocks = sampler.stop()

    # ---- end-to-end through the public API with HOST buffers (pinned), copies inside the timed region ----
    n_codes = 32 * (T // 32 + (T + 31) // 32)

    def pinned_out(b, t):
        return (torch.empty(b, 80, t, dtype=torch.float32).pin_memory(), torch.empty(b, 80, t, dtype=torch.float32).pin_memory(),
                torch.empty(b, 32 * (t // 32 + (t + 31) // 32), dtype=torch.float32).pin_memory())

    def e2e_pipeline(mdl, host_in, host_out, steps):
        """HostPipeline (autovc_b200/voice_converter.py): H2D of step i+1 and D2H of step i-1 ride two copy streams while
        step i computes; every step copies its own inputs in and all three outputs out inside the timed region."""
        pipe = HostPipeline(mdl, dev)
        ins, outs = [host_in] * steps, [host_out] * steps
        pipe.run(ins[:max(2, min(steps, 6))], outs[:max(2, min(steps, 6))])   # untimed: fills the allocator's pools of the streams
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        flush.zero_()
        barrier()
        e0.record()
        pipe.run(ins, outs)
        e1.record()
        barrier()
        return e0.elapsed_time(e1)

    host_out = pinned_out(B, T)
    e2e_ms = e2e_pipeline(model, (x_h, co_h, ct_h), host_out, K)
    h2d = x_h.numel() * 4 + co_h.numel() * 4 + ct_h.numel() * 4
    d2h = sum(t.numel() * 4 for t in host_out)

    # ---- per-phase device time (CUDA events inside the library, same stream) for the roofline ----
    def phases(mdl, inputs, reps):
        e = mdl._engine(dev)
        e.set_profiling(True)
        acc = {}
        for _ in range(reps):
            flush.zero_()
            mdl(*inputs)
            for k, v in e.profile().items():
                acc[k] = acc.get(k, 0.0) + v / reps
        e.set_profiling(False)
        return {k: round(v, 4) for k, v in acc.items()}

    phase_ms = phases(model, (x, c_org, c_trg), min(K, 5))

    # ---- p50 latency of a single 128-frame utterance (B=1) ----
    def p50(fn, warm, reps):
        for _ in range(warm):
            fn()
        lat = []
        for _ in range(reps):
            s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            s.record(); fn(); e.record(); e.synchronize()
            lat.append(s.elapsed_time(e))
        return statistics.median(lat)

    x1, co1, ct1 = x[:1].contiguous(), c_org[:1].contiguous(), c_trg[:1].contiguous()
    lat_p50 = p50(lambda: model(x1, co1, ct1), 3, 50)

    # per-rank record (N > 1): every rank's own step time and SM clock under load, gathered outside the timed regions
    per_rank = None
    if dist is not None:
        mine = torch.tensor([dev_ms / K, e2e_ms / K, float(clocks.get("sm_mhz") or 0.0)], dtype=torch.float64, device=dev)
        allr = [torch.empty_like(mine) for _ in range(world)]
        dist.all_gather(allr, mine)
        per_rank = {"ms_per_step": [round(float(t[0]), 4) for t in allr], "e2e_ms_per_step": [round(float(t[1]), 4) for t in allr],
                    "sm_mhz": [float(t[2]) for t in allr]}
    dev_ms, e2e_ms = max_over_ranks(dev_ms, e2e_ms)

    # ---- the other BASELINE.json configs, same process, a few steps each (device events, max over ranks) ----
    configs = {}
    if not args.no_configs and args.precision == "bf16":
        # latency split at B = 1
        configs["b1"] = {"workload": "B=1 x T=128 (one utterance)", "p50_ms": lat_p50, "phase_ms": phases(model, (x1, co1, ct1), 5)}

        # C3: long-utterance LSTM stress, B = 64 x T = 1024 (4097 dependent recurrence steps)
        B3, T3, K3 = 64, 1024, 5
        x3_h, co3_h, ct3_h = (t.pin_memory() for t in synth_inputs(B3, T3, seed=77 + rank))
        x3, co3, ct3 = x3_h.to(dev), co3_h.to(dev), ct3_h.to(dev)
        ms3 = timed_steps(lambda: model(x3, co3, ct3), 2, K3)
        ph3 = phases(model, (x3, co3, ct3), 2)
        e3 = e2e_pipeline(model, (x3_h, co3_h, ct3_h), pinned_out(B3, T3), K3)
        ms3, e3 = max_over_ranks(ms3, e3)
        configs["c3"] = {"workload": f"config3: B={B3} x T={T3} per GPU", "steps": K3, "ms": ms3 / K3,
                         "value": world * B3 * T3 * K3 / (ms3 * 1e-3), "unit": "frames/s", "phase_ms": ph3,
                         "e2e": {"value": world * B3 * T3 * K3 / (e3 * 1e-3), "unit": "frames/s"}}
        del x3, co3, ct3, x3_h

        # C5: training-shape forward pair (reference model.py:306 then :201), 256 utterances per GPU
        def pair(mdl):
            _mel, post, _codes = mdl(x, c_org, c_org)
            mdl(post, c_org, None)
        K5 = 5
        ms5 = timed_steps(lambda: pair(model), 2, K5)
        tm = autovc_b200.AutoEncoder(device=dev, verbose=False, precision=args.precision, batch_stat_bn=True)
        tm.load_state_dict(sd)
        tm.train()
        ms5t = timed_steps(lambda: pair(tm), 2, K5)
        e5 = e2e_pipeline(model, (x_h, co_h, co_h), host_out, K5)     # pass 1 from / to host buffers (the pair's second pass stays on the device)
        ms5, ms5t, e5 = max_over_ranks(ms5, ms5t, e5)
        configs["c5"] = {"workload": f"config5: forward(x,c,c) + forward(mel_post,c,None), B={B} x T={T} per GPU (2048 over 8 GPUs)",
                         "steps": K5, "ms": ms5 / K5, "value": world * B * T * K5 / (ms5 * 1e-3), "unit": "frames/s",
                         "batch_stat_bn": {"ms": ms5t / K5, "value": world * B * T * K5 / (ms5t * 1e-3)},
                         "e2e": {"value": world * B * T * K5 / (e5 * 1e-3), "unit": "frames/s", "note": "pass 1 through HostPipeline"}}
        del tm

        # C4: VoiceConverter.convert_multiple, 4096 utterances x T=128 over all ranks, 64 speakers, the d-vector encoder
        # once per speaker; sources are rows of pinned HOST memory, results come back to pinned host memory
        from autovc_b200.synth import synth_speaker_encoder
        N4, S4, T4 = 4096, 64, 128
        per = (N4 + world - 1) // world
        lo, hi = min(N4, rank * per), min(N4, (rank + 1) * per)
        se_sd, _parts = synth_speaker_encoder(5)
        se = autovc_b200.SpeakerEncoder(device=dev, verbose=False)
        se.load_state_dict(se_sd)
        g = torch.Generator().manual_seed(4242)
        partials = [torch.randn(8, 160, 40, generator=g) for _ in range(S4)]          # [8,160,40] N(0,1)-derived partials per speaker
        slab = torch.rand(hi - lo, 80, T4, generator=torch.Generator().manual_seed(99 + rank)).pin_memory()
        dummy = torch.zeros(80, 1)
        # every rank builds the same work list; only its own block carries real mels (the other ranks' rows are never touched)
        srcs = [Utterance(slab[i - lo] if lo <= i < hi else dummy, f"spk{i * S4 // N4}", partials[i * S4 // N4]) for i in range(N4)]
        trgs = [Utterance(dummy, f"spk{(i * S4 // N4 + 1) % S4}", partials[(i * S4 // N4 + 1) % S4]) for i in range(N4)]
        vc = VoiceConverter(model, embed_speaker=se.as_embed_fn(), device=dev, max_batch=B)
        mine_keys = [u.speaker for u in srcs[lo:hi]] + [u.speaker for u in trgs[lo:hi]]
        mine_data = [u.speaker_data for u in srcs[lo:hi]] + [u.speaker_data for u in trgs[lo:hi]]
        # untimed passes first: the speaker encoder's workspace / plans for this batch of partials, and torch's caching HOST allocator,
        # which needs a few rounds before the 168 MB pinned result block of a pass is recycled instead of cudaHostAlloc'ed (~100 ms)
        vc.learn_speakers(mine_keys, mine_data, embed_many=se.embed_speakers)
        out4 = {(80, T4): torch.empty(hi - lo, 80, T4).pin_memory()}      # the serving loop's own result buffer, reused from call to call
        c4_args = dict(match_method="align", distributed=world > 1, gather=False, output="host", host_buffers=out4)
        for _ in range(2):
            vc.convert_multiple(srcs, trgs, **c4_args)
        vc.speakers.clear()
        vc.embed_calls = 0
        barrier()
        a0, a1, s1, s2 = (torch.cuda.Event(enable_timing=True) for _ in range(4))
        a0.record()
        n_emb = vc.learn_speakers(mine_keys, mine_data, embed_many=se.embed_speakers)     # ONE batched call, once per speaker
        a1.record()
        barrier()
        flush.zero_()
        barrier()
        s1.record()
        res = vc.convert_multiple(srcs, trgs, **c4_args)
        s2.record()
        barrier()
        assert sum(r is not None for r in res) == hi - lo and vc.embed_calls == n_emb
        se_ms, c4_ms = max_over_ranks(a0.elapsed_time(a1), s1.elapsed_time(s2))
        configs["c4"] = {"workload": f"config4: VoiceConverter.convert_multiple, {N4} utterances x T={T4} in total ({hi - lo} per rank), {S4} speakers, "
                                     "pinned host mels in, pinned host mel_outputs_postnet out (the call is end to end by construction)",
                         "ms": c4_ms, "value": N4 * T4 / (c4_ms * 1e-3), "unit": "frames/s",
                         "e2e": {"value": N4 * T4 / (c4_ms * 1e-3), "unit": "frames/s",
                                 "h2d_bytes": (hi - lo) * (80 * T4 + 512) * 4, "d2h_bytes": (hi - lo) * 80 * T4 * 4},
                         "speaker_encoder": {"ms": se_ms, "speakers_per_rank": n_emb, "partials_per_speaker": 8,
                                             "note": "all speakers of a rank in one batched SpeakerEncoder.embed_speakers call, outside the convert timing"}}
        del slab, srcs, trgs, res, out4

    if rank == 0:
        peaks, peak_src = load_peaks()
        frames_total = world * B * T * K
        value = frames_total / (dev_ms * 1e-3)
        e2e_value = frames_total / (e2e_ms * 1e-3)
        dom = max((k for k in phase_ms if k in MAC_PER_FRAME), key=lambda k: phase_ms[k])
        dom_flops = 2.0 * MAC_PER_FRAME[dom] * B * T
        achieved = dom_flops / (phase_ms[dom] * 1e-3) / 1e12 if phase_ms[dom] > 0 else 0.0
        peak = float(peaks.get("bf16_tflops_sustained", peaks.get("bf16_tflops")))
        burst = float(peaks.get("bf16_tflops", peak))
        # DRAM bytes of the dominant phase's launches from the committed ncu --set full capture (same config), or null
        traffic = None
        for name in ("r2_ncu_traffic.json", "r1_ncu_traffic.json"):
            try:
                tj = json.load(open(os.path.join(ROOT, "profiles", name)))
                if (B, T) == (256, 128) and args.precision == "bf16":
                    traffic = tj["bytes_per_step"].get(dom)
                break
            except Exception:
                continue
        line = {
            "metric": "generator_fwd_mel_frames_per_s", "value": value, "unit": "frames/s", "n_gpus": world,
            "steps": K, "warmup": W, "ms_per_step": dev_ms / K, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "bf16" if args.precision == "bf16" else "f32", "data": "synthetic",
            "config": {"workload": f"config2: B={B} x T={T} x 80 mel per GPU, dims 32/256/512/32, seeded random-init weights",
                       "l2": "256 MB flush written between timed iterations (excluded from the timing)",
                       "parallelism": f"replicas x{world}, no data-path collective",
                       "e2e": "HostPipeline: pinned host buffers, per-step H2D and D2H on two copy streams overlapped with the previous/next step's compute",
                       "launch": "the forward's kernel launches (26 at config 2) are replayed as one CUDA graph built by the library on the second call with the same pointers"},
            "realtime_x": value / REALTIME_FRAMES_PER_S,
            "whole_path_tensor_frac": value / world * FLOP_PER_FRAME / (peak * 1e12),
            "whole_path_tensor_frac_burst": value / world * FLOP_PER_FRAME / (burst * 1e12),
            "latency_p50_ms_b1_t128": lat_p50,
            "e2e": {"value": e2e_value, "unit": "frames/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h},
            "gpu_launches": int(launches_per_step * K),
            "clocks": clocks,
            "phase_ms": phase_ms,
            "roofline": {"bound": "tensor", "kernel": dom, "achieved": achieved, "peak": peak, "unit": "TFLOP/s",
                         "frac": achieved / peak, "frac_of_burst_peak": achieved / burst, "traffic": traffic,
                         "peak_source": f"{peak_src} bf16_tflops_sustained ({peak}); burst {burst}",
                         "algorithmic_flop_per_launch_group": dom_flops},
            "wall_s_timed_region": wall,
        }
        if per_rank:
            line["per_rank"] = per_rank
        if configs:
            line["configs"] = configs
        if world == 1 and not args.no_configs:
            line["gpu_eager_baseline"] = gpu_eager_baseline(dev, B, T, x, c_org, c_trg)
        if not args.no_cpu_baseline and world == 1:     # the CPU leg is an N=1 measurement (rank 0 would share the host with the other ranks)
            r = time_cpu_reference(B, T, 3, 1, budget_s=40.0)
            line["cpu_baseline"] = {k: r[k] for k in ("value", "unit", "cores", "kind", "sample")}
        print(json.dumps(line))
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
