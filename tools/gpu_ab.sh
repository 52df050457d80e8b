#!/bin/bash
# A/B an env setting: parity errors (printed) + bench for each config string
mkdir -p gpurun_out
for cfg in "$@"; do echo "=== $cfg"
env $cfg timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -q -s -k "bf16_mode_matches" 2>&1 | grep -E "bf16 mel|passed|failed"
env $cfg timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/b_ab.log 2>&1; python -c "
import json;d=json.loads(open('gpurun_out/b_ab.log').read().strip().splitlines()[-1]);print(d['value'],d['ms_per_step'],d['phase_ms'])" || tail -5 gpurun_out/b_ab.log
done
