#!/bin/bash
# config-3 phase split for prebuilt libraries on the same box: tools/gpu_c3_ab.sh libA.so libB.so ...
mkdir -p gpurun_out
cp autovc_b200/lib/libautovc_b200.so /tmp/orig.so
for rep in 1 2 3; do for lib in "$@"; do
cp $lib autovc_b200/lib/libautovc_b200.so
timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/b_c3.log 2>&1; python -c "
import json,sys;d=json.loads(open('gpurun_out/b_c3.log').read().strip().splitlines()[-1]);c=d['configs']['c3']
print(sys.argv[1], round(d['ms_per_step'],4), 'c3', round(c['ms'],3), c['phase_ms'])" $lib || tail -5 gpurun_out/b_c3.log
done; done
cp /tmp/orig.so autovc_b200/lib/libautovc_b200.so
