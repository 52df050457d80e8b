#!/bin/bash
# LSTM kernel iteration: bf16 parity -> per-step timeline -> short bench.  Extra env (e.g. AVC_LSTM_SPEC=2) is passed through.
mkdir -p gpurun_out
echo "== bf16 parity"; timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -q -k "bf16 or alias or training" -x > gpurun_out/t_bf16.log 2>&1; echo "exit $?"; tail -15 gpurun_out/t_bf16.log
echo "== trace"; AVC_LSTM_TRACE=1 PW=1 PK=1 timeout 300 python tools/profile_step.py 2>&1 | grep -A9 "trace" | head -60
echo "== bench"; timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/b_bf16.log 2>&1; echo "exit $?"; python -c "
import json;d=json.loads(open('gpurun_out/b_bf16.log').read().strip().splitlines()[-1]);print(d['value'],d['e2e']['value'],d['ms_per_step'],d['phase_ms'],d['latency_p50_ms_b1_t128'])" || tail -20 gpurun_out/b_bf16.log
