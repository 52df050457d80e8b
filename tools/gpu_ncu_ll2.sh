#!/bin/bash
mkdir -p gpurun_out
PW=1 PK=1 python tools/profile_step.py > gpurun_out/plain.log 2>&1 &&
PW=1 PK=1 ncu --set full --clock-control none --import-source on -k regex:ll2_lstm -s 2 -c 1 -f -o gpurun_out/prof_ll2 python tools/profile_step.py > gpurun_out/ncu_ll2.log 2>&1
echo "ncu exit $?"; tail -3 gpurun_out/ncu_ll2.log; ls -la gpurun_out/prof_ll2.ncu-rep
