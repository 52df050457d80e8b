#!/bin/bash
# per-step timeline of the LSTM kernels for a list of env settings: tools/gpu_trace.sh "A=1 B=2" "C=3" ...
mkdir -p gpurun_out
for cfg in "$@"; do echo "--- $cfg"; env $cfg AVC_LSTM_TRACE=1 PW=0 PK=1 timeout 300 python tools/profile_step.py 2>&1 | grep -A9 "trace" | head -22; done
