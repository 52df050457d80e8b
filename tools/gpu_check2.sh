#!/bin/bash
mkdir -p gpurun_out
echo "== bf16 parity"; timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -q -k "bf16 or alias" -x > gpurun_out/t_bf16.log 2>&1; echo "exit $?"; tail -3 gpurun_out/t_bf16.log
echo "== bench bf16"; timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/b_bf16.log 2>&1; echo "exit $?"; python -c "
import json;d=json.loads(open('gpurun_out/b_bf16.log').read().strip().splitlines()[-1]);print(d['value'],d['e2e']['value'],d['ms_per_step'],d['phase_ms'],d['latency_p50_ms_b1_t128'])"
