#!/bin/bash
mkdir -p gpurun_out
for skb in 2 4; do echo "--- skb=$skb"; AVC_LSTM_SKB=$skb AVC_LSTM_TRACE=1 PW=1 PK=1 python tools/profile_step.py 2>&1 | grep -A4 "H=1024" | tail -5; AVC_LSTM_SKB=$skb AVC_LSTM_TRACE=1 PW=1 PK=1 python tools/profile_step.py 2>&1 | grep -A3 "H=512" | head -4; done
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -q -k "bf16" -x 2>&1 | tail -3
timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/b_bf16.log 2>&1; echo "exit $?"; python -c "
import json;d=json.loads(open('gpurun_out/b_bf16.log').read().strip().splitlines()[-1]);print(d['value'],d['ms_per_step'],d['phase_ms'],d['latency_p50_ms_b1_t128'])"
