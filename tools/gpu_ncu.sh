#!/bin/bash
# ncu captures of config 2 (one forward): launch list, then --set full of the recurrence and GEMM kernels.
# The cooperative cluster launch falls back to the checked plain cluster launch under a profiler (launch.h::launch_persistent).
mkdir -p gpurun_out
timeout 300 python tools/profile_step.py > gpurun_out/plain.log 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -s 52 -c 26 --csv --log-file gpurun_out/launches.csv python tools/profile_step.py > gpurun_out/ncu1.log 2>&1
echo "launch list exit $?"; tail -2 gpurun_out/plain.log
PW=1 timeout 900 ncu --set full --clock-control none --import-source on -k regex:ll_lstm_kernel -s 3 -c 3 -o gpurun_out/prof_lstm -f python tools/profile_step.py > gpurun_out/ncu2.log 2>&1
echo "lstm full exit $?"; tail -3 gpurun_out/ncu2.log
PW=1 timeout 900 ncu --set full --clock-control none --import-source on -k regex:tc_gemm_kernel -s 18 -c 18 -o gpurun_out/prof_gemm -f python tools/profile_step.py > gpurun_out/ncu3.log 2>&1
echo "gemm full exit $?"; tail -3 gpurun_out/ncu3.log
ls -la gpurun_out
timeout 900 python bench.py > gpurun_out/b_bf16.log 2> gpurun_out/b_bf16.err; echo "bench exit $?"
