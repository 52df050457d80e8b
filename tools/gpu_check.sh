#!/bin/bash
# One gpurun call: fp32 parity -> tensor-core GEMM unit tests -> bf16 parity -> short benches.  Logs in gpurun_out/.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/gpu.txt 2>&1
echo "== fp32 parity"; timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -q -k "fp32 or training" -x > gpurun_out/t_fp32.log 2>&1; echo "exit $?"; tail -5 gpurun_out/t_fp32.log
echo "== tc gemm unit"; timeout 600 python -m pytest tests/test_gpu_tc_gemm.py -m gpu -q > gpurun_out/t_tc.log 2>&1; echo "exit $?"; tail -25 gpurun_out/t_tc.log
echo "== bf16 parity"; timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -q -k "bf16" -s > gpurun_out/t_bf16.log 2>&1; echo "exit $?"; tail -25 gpurun_out/t_bf16.log
echo "== bench fp32"; timeout 600 python bench.py --precision fp32 --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/b_fp32.log 2>&1; echo "exit $?"; tail -3 gpurun_out/b_fp32.log
echo "== bench bf16"; timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/b_bf16.log 2>&1; echo "exit $?"; tail -3 gpurun_out/b_bf16.log
echo "== bench bf16 forced simt gemm"; AVC_FORCE_SIMT=1 timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/b_bf16_simt.log 2>&1; echo "exit $?"; tail -3 gpurun_out/b_bf16_simt.log
