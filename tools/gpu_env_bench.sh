#!/bin/bash
# bench (phase times) for a list of env settings: tools/gpu_env_bench.sh "A=1" "B=2" ...
mkdir -p gpurun_out
for cfg in "$@"; do echo "=== $cfg"
env $cfg timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/b_env.log 2>&1; python -c "
import json;d=json.loads(open('gpurun_out/b_env.log').read().strip().splitlines()[-1]);print(d['value'],d['ms_per_step'],d['phase_ms'])" || tail -5 gpurun_out/b_env.log
done
