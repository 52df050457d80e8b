"""Vocoder measurement: samples/s of the per-sample loop on cuda:0, the end-to-end `generate_batch`
(host mels in, host waveforms out), and the numpy oracle on the host cores beside it.  One JSON line.

    python tools/bench_vocoder.py [--rows 512] [--steps 2000] [--utts 128 --frames 128]

Algorithmic work per sample and row (the reference's dense formulation, model.py:277-294):
I 113x512 + GRU1 2x1536x512 + GRU2 1536x(544+512) + fc1 544x512 + fc2 544x512 + fc3 512x30 = 3,825,152 MAC = 7,650,304 FLOP.
"""
import argparse, json, os, sys, time
import numpy as np
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from autovc_b200.vocoder import WaveRNN, fold_params
from autovc_b200.synth import synth_wavernn_state_dict, synth_uniforms

FLOP_PER_SAMPLE = 2 * 3_825_152
BYTES_PER_ROW_STEP = 53 * 16384 / 128      # operand blocks every CTA brings in per tile and step, per row


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rows", type=int, default=512)
    ap.add_argument("--steps", type=int, default=2000)
    ap.add_argument("--utts", type=int, default=128)
    ap.add_argument("--frames", type=int, default=128)
    ap.add_argument("--cpu-steps", type=int, default=30)
    ap.add_argument("--no-e2e", action="store_true")
    a = ap.parse_args()
    sd = synth_wavernn_state_dict(11)
    voc = WaveRNN(device="cuda")
    voc.load_state_dict(sd)
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")))
    except Exception:
        pass
    mel = torch.rand(4, 80, 64, device="cuda")
    m_up, aux = voc.condition(mel)
    out = {"metric": "vocoder samples/s (WaveRNN generate loop)", "unit": "samples/s", "dtype": "bf16", "data": "synthetic", "loop": []}
    for rows in sorted({1, 128, 256, a.rows}):
        ru = (np.arange(rows) % 4).astype(np.int32)
        rs = ((np.arange(rows) * 97) % (60 * 275)).astype(np.int32)
        voc.sample_rows(m_up, aux, ru, rs, 100)
        torch.cuda.synchronize()
        ts = []
        for _ in range(3):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); s = voc.sample_rows(m_up, aux, ru, rs, a.steps); e1.record(); torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1))
        ms = float(np.median(ts))
        sps = rows * a.steps / ms * 1e3
        out["loop"].append({"rows": rows, "steps": a.steps, "ms": round(ms, 2), "us_per_step": round(ms / a.steps * 1e3, 2),
                            "samples_per_s": round(sps), "x_realtime_22050": round(sps / 22050, 1),
                            "tflops": round(sps * FLOP_PER_SAMPLE / 1e12, 2),
                            "l2_to_sm_GBps": round(sps * BYTES_PER_ROW_STEP * 128 / 1e9, 1), "launches": voc.last_launch_count})
    out["value"] = out["loop"][-1]["samples_per_s"]
    out["config"] = {"workload": f"{a.rows} rows x {a.steps} steps, reference hyper-parameters (hparams.py:92-107), random weights"}
    if not a.no_e2e:
        # end to end: U utterances of T frames, host mels -> conditioning -> folds -> loop -> host waveforms (cross-fade on the host)
        mels = [torch.rand(1, 80, a.frames) for _ in range(a.utts)]
        voc.generate_batch(mels[:2])
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        waves = voc.generate_batch(mels)
        dt = time.perf_counter() - t0
        n = sum(len(w) for w in waves)
        out["e2e"] = {"value": round(n / dt), "unit": "samples/s", "utterances": a.utts, "frames": a.frames, "rows": voc.last_rows,
                      "steps": voc.last_steps, "seconds": round(dt, 3), "audio_seconds": round(n / 22050, 1),
                      "x_realtime_22050": round(n / dt / 22050, 1), "h2d_bytes": a.utts * 80 * a.frames * 4, "d2h_bytes": voc.last_rows * voc.last_steps * 4}
    # the oracle (numpy, float32) on the host cores: rows x cpu_steps of the same loop
    from oracle import wavernn_oracle as wo
    rows = 128
    mm = np.random.default_rng(0).uniform(0, 1, (rows, a.cpu_steps, 80)).astype(np.float32)
    aa = np.random.default_rng(1).standard_normal((rows, a.cpu_steps, 128)).astype(np.float32)
    U = synth_uniforms(a.cpu_steps, rows, 3)
    wo.generate_rows(sd, mm[:, :3], aa[:, :3], U[:3])
    t0 = time.perf_counter()
    wo.generate_rows(sd, mm, aa, U)
    dt = time.perf_counter() - t0
    out["cpu_baseline"] = {"value": round(rows * a.cpu_steps / dt), "unit": "samples/s", "cores": os.cpu_count(), "kind": "port",
                           "sample": f"{rows} rows x {a.cpu_steps} steps of the loop, numpy float32 (BLAS threads: all)"}
    print(json.dumps(out))


if __name__ == "__main__":
    main()
