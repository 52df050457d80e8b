#!/bin/bash
# tc gemm unit tests + bf16 parity + short bench
mkdir -p gpurun_out
echo "== tc gemm unit"; timeout 600 python -m pytest tests/test_gpu_tc_gemm.py -m gpu -q -x 2>&1 | tail -3
echo "== bf16 parity"; timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -q -k "bf16 or alias or training" -x 2>&1 | tail -3
echo "== bench"; timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/b_bf16.log 2>&1; echo "exit $?"; python -c "
import json;d=json.loads(open('gpurun_out/b_bf16.log').read().strip().splitlines()[-1]);print(d['value'],d['e2e']['value'],d['ms_per_step'],d['phase_ms'],d['latency_p50_ms_b1_t128'])" || tail -20 gpurun_out/b_bf16.log
