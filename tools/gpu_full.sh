#!/bin/bash
# full GPU suite + smoke + bench line
mkdir -p gpurun_out
echo "== all gpu tests"; timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/t_all.log 2>&1; echo "exit $?"; tail -6 gpurun_out/t_all.log
echo "== smoke"; timeout 300 python __graft_entry__.py smoke 2>&1 | tail -4
echo "== bench"; timeout 900 python bench.py > gpurun_out/b_full.log 2> gpurun_out/b_full.err; echo "exit $?"; tail -3 gpurun_out/b_full.err; python - <<'PY'
import json
d=json.loads(open('gpurun_out/b_full.log').read().strip().splitlines()[-1])
print('value', round(d['value']), 'ms', round(d['ms_per_step'],4), 'e2e', round(d['e2e']['value']), 'p50', round(d['latency_p50_ms_b1_t128'],3))
print('phase', d['phase_ms']); print('roofline', {k:(round(v,4) if isinstance(v,float) else v) for k,v in d['roofline'].items() if k in ('kernel','achieved','frac','frac_of_burst_peak')})
for k,v in d.get('configs',{}).items(): print(k, {a:(round(b,3) if isinstance(b,float) else b) for a,b in v.items() if a in ('ms','value','p50_ms','phase_ms','e2e','batch_stat_bn','speaker_encoder')})
print('eager', d.get('gpu_eager_baseline')); print('cpu', d.get('cpu_baseline',{}).get('value'))
PY
