#!/bin/bash
# quick: lstm-relevant parity + 2 bench runs (default build)
mkdir -p gpurun_out
echo "== parity subset"; timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_speaker_encoder.py -m gpu -q -x -k "not stress" 2>&1 | tail -3
for rep in 1 2; do
  timeout 600 python bench.py --no-configs --no-cpu-baseline --steps 20 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print(round(d['ms_per_step'],4), d['phase_ms'], 'p50', round(d['latency_p50_ms_b1_t128'],3), 'e2e', round(d['e2e']['value']/1e6,2))
    elif 'rror' in l: print(l.strip()[:300])
"
done
