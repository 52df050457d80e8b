#!/bin/bash
for d in 0 1 2 3; do echo "=== AVC_LSTM_DBG=$d"; env AVC_LSTM_DBG=$d timeout 600 python bench.py --no-configs --no-cpu-baseline --steps 20 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print(round(d['ms_per_step'],4), d['phase_ms'])
    elif 'rror' in l: print(l.strip()[:300])
"; done
