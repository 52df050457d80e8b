#!/bin/bash
# ncu captures of config 2 (one forward): launch list, then --set full of the recurrence and GEMM kernels.
# AVC_LSTM_NOCOOP=1: ncu fails cooperative cluster launches ("LaunchFailed"); the plain cluster launch is the same grid.
mkdir -p gpurun_out
export AVC_LSTM_NOCOOP=1
timeout 300 python tools/profile_step.py > gpurun_out/plain.log 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -s 48 -c 24 --csv --log-file gpurun_out/launches.csv python tools/profile_step.py > gpurun_out/ncu1.log 2>&1
echo "launch list exit $?"; tail -2 gpurun_out/plain.log
PW=1 timeout 900 ncu --set full --clock-control none --import-source on -k regex:ll_lstm_kernel -s 3 -c 3 -o gpurun_out/prof_lstm -f python tools/profile_step.py > gpurun_out/ncu2.log 2>&1
echo "lstm full exit $?"; tail -3 gpurun_out/ncu2.log
PW=1 timeout 900 ncu --set full --clock-control none --import-source on -k regex:tc_gemm_kernel -s 19 -c 19 -o gpurun_out/prof_gemm -f python tools/profile_step.py > gpurun_out/ncu3.log 2>&1
echo "gemm full exit $?"; tail -3 gpurun_out/ncu3.log
ls -la gpurun_out
