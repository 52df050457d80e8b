#!/bin/bash
# A/B of environment settings over the whole bench (configs included): tools/gpu_ab_env.sh "A=1" "B=2" ...
mkdir -p gpurun_out
echo "== parity subset"; timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_speaker_encoder.py -m gpu -q -x 2>&1 | tail -3
for cfg in "$@"; do
  echo "--- $cfg"; env $cfg timeout 900 python bench.py --no-cpu-baseline --steps 20 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print(round(d['ms_per_step'],4), d['phase_ms'], 'p50', round(d['latency_p50_ms_b1_t128'],3), 'e2e', round(d['e2e']['value']/1e6,2))
        c=d.get('configs',{})
        print('  b1', c.get('b1',{}).get('phase_ms')); print('  c3', round(c['c3']['ms'],3), c['c3']['phase_ms']); print('  c5', round(c['c5']['ms'],3), round(c['c5']['batch_stat_bn']['ms'],3), ' c4', round(c['c4']['ms'],2), round(c['c4']['value']/1e6,2), 'se', round(c['c4']['speaker_encoder']['ms'],2))
    elif 'rror' in l: print(l.strip()[:300])
"
done
