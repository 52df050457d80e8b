#!/bin/bash
mkdir -p gpurun_out
echo "== bf16 parity"; timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -q -k "bf16" -s -x > gpurun_out/t_bf16.log 2>&1; echo "exit $?"; tail -25 gpurun_out/t_bf16.log
echo "== bench bf16"; timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/b_bf16.log 2>&1; echo "exit $?"; tail -3 gpurun_out/b_bf16.log
