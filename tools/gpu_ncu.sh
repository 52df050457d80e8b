#!/bin/bash
mkdir -p gpurun_out
python tools/profile_step.py > gpurun_out/plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -s 48 -c 48 --csv --log-file gpurun_out/launches.csv python tools/profile_step.py > gpurun_out/ncu1.log 2>&1
echo "launch list exit $?"; tail -2 gpurun_out/plain.log
PW=1 python tools/profile_step.py > gpurun_out/plain2.log 2>&1 && \
PW=1 ncu --set full --clock-control none --import-source on -k regex:tc_lstm_kernel -s 3 -c 3 -o gpurun_out/prof_lstm -f python tools/profile_step.py > gpurun_out/ncu2.log 2>&1
echo "lstm full exit $?"; tail -3 gpurun_out/ncu2.log
PW=1 ncu --set full --clock-control none --import-source on -k regex:tc_gemm_kernel -s 19 -c 4 -o gpurun_out/prof_gemm -f python tools/profile_step.py > gpurun_out/ncu3.log 2>&1
echo "gemm full exit $?"; tail -3 gpurun_out/ncu3.log
ls -la gpurun_out
