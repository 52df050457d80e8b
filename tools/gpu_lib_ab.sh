#!/bin/bash
# A/B prebuilt libraries on the same box: tools/gpu_lib_ab.sh libA.so libB.so ...   (3 bench runs each, interleaved; AB_CONFIGS=1: with c3 / b1)
mkdir -p gpurun_out
cp autovc_b200/lib/libautovc_b200.so /tmp/orig.so
for rep in 1 2 3; do for lib in "$@"; do
cp $lib autovc_b200/lib/libautovc_b200.so
timeout 600 python bench.py --steps 20 --warmup 3 --no-cpu-baseline $( [ -z "$AB_CONFIGS" ] && echo --no-configs ) > gpurun_out/b_lib.log 2>&1; python -c "
import json,sys;d=json.loads(open('gpurun_out/b_lib.log').read().strip().splitlines()[-1]);c=d.get('configs',{})
print(sys.argv[1], round(d['ms_per_step'],4),d['phase_ms'],'p50',round(d['latency_p50_ms_b1_t128'],3), 'c3', round(c['c3']['ms'],3) if c else '', 'c4', round(c['c4']['ms'],2) if c else '')" $lib || tail -5 gpurun_out/b_lib.log
done; done
cp /tmp/orig.so autovc_b200/lib/libautovc_b200.so
