#!/bin/bash
# A/B two prebuilt libraries on the same box: tools/gpu_lib_ab.sh libA.so libB.so   (3 bench runs each, interleaved)
mkdir -p gpurun_out
cp autovc_b200/lib/libautovc_b200.so /tmp/orig.so
for rep in 1 2 3; do for lib in "$@"; do
cp $lib autovc_b200/lib/libautovc_b200.so
timeout 600 python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-configs > gpurun_out/b_lib.log 2>&1; python -c "
import json,sys;d=json.loads(open('gpurun_out/b_lib.log').read().strip().splitlines()[-1]);print(sys.argv[1], round(d['ms_per_step'],4),d['phase_ms'],'p50',round(d['latency_p50_ms_b1_t128'],3))" $lib || tail -5 gpurun_out/b_lib.log
done; done
cp /tmp/orig.so autovc_b200/lib/libautovc_b200.so
