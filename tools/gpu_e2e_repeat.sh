#!/bin/bash
# e2e stability: N default bench runs, value / e2e per run
mkdir -p gpurun_out
for i in $(seq 1 ${1:-4}); do timeout 600 python bench.py --no-cpu-baseline > gpurun_out/b_rep.log 2>&1; python -c "
import json;d=json.loads(open('gpurun_out/b_rep.log').read().strip().splitlines()[-1]);print(round(d['value']),round(d['e2e']['value']),round(d['ms_per_step'],4),d['steps'])" || tail -3 gpurun_out/b_rep.log; done
