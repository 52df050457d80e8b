#!/bin/bash
mkdir -p gpurun_out
echo "== all gpu tests"; timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/t_all.log 2>&1; echo "exit $?"; tail -4 gpurun_out/t_all.log
echo "== smoke"; timeout 300 python __graft_entry__.py smoke 2>&1 | tail -3
echo "== bench"; timeout 600 python bench.py > gpurun_out/b_bf16.log 2>&1; echo "exit $?"; python -c "
import json;d=json.loads(open('gpurun_out/b_bf16.log').read().strip().splitlines()[-1]);print(d['value'],d['e2e']['value'],d['ms_per_step'],d['phase_ms'],d['latency_p50_ms_b1_t128'],d['cpu_baseline']['value'], d['clocks'])"
echo "== reference arm"; timeout 600 python bench.py --impl reference --steps 5 --warmup 1 | cut -c1-200
