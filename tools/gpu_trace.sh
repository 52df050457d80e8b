#!/bin/bash
# per-step timeline of the recurrence kernels (clock64 stamps of CTA (0,0), steps 8..15): tools/gpu_trace.sh ["ENV=1 ..."]
env $1 AVC_LSTM_TRACE=1 PW=1 PK=1 PB=${PB:-256} PT=${PT:-128} timeout 300 python tools/profile_step.py 2>&1 | grep -A17 "ll_lstm trace" | cut -c1-220
