#!/bin/bash
# A/B env settings: bench (3 repeats each, interleaved) + B=1 latency; parity of the last config
mkdir -p gpurun_out
for rep in 1 2; do for cfg in "$@"; do
env $cfg timeout 600 python bench.py --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/b_ab.log 2>&1; python -c "
import json;d=json.loads(open('gpurun_out/b_ab.log').read().strip().splitlines()[-1]);print('$cfg', round(d['ms_per_step'],4),d['phase_ms'],'p50',round(d['latency_p50_ms_b1_t128'],4))" || tail -5 gpurun_out/b_ab.log
done; done
for cfg in "$@"; do echo "== parity $cfg"; env $cfg timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -q -k "bf16_mode_matches or beyond_one" 2>&1 | tail -1; done
