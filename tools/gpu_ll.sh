#!/bin/bash
# recurrence kernel check: bf16 parity + speaker encoder (H=256) + row independence, then step trace, then bench
mkdir -p gpurun_out
echo "== parity"; timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_speaker_encoder.py -m gpu -q -x -k "bf16 or alias or training or independent" 2>&1 | tail -5
echo "== trace"; AVC_LSTM_TRACE=1 PW=0 PK=1 timeout 300 python tools/profile_step.py 2>&1 | grep -A9 "trace" | head -32
echo "== bench"; timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/b_bf16.log 2>&1; echo "exit $?"; python -c "
import json;d=json.loads(open('gpurun_out/b_bf16.log').read().strip().splitlines()[-1]);print(d['value'],d['e2e']['value'],d['ms_per_step'],d['phase_ms'],d['latency_p50_ms_b1_t128'])" || tail -20 gpurun_out/b_bf16.log
