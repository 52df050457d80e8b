#!/bin/bash
# one ncu --set full capture (with source counters) of a recurrence kernel launch: tools/gpu_ncu_ll.sh [skip-launches]
mkdir -p gpurun_out
PW=1 PK=1 python tools/profile_step.py > gpurun_out/plain.log 2>&1 &&
PW=1 PK=1 ncu --set full --clock-control none --import-source on -k regex:ll_lstm -s ${1:-2} -c 1 -f -o gpurun_out/prof_ll python tools/profile_step.py > gpurun_out/ncu_ll.log 2>&1
echo "ncu exit $?"; tail -3 gpurun_out/ncu_ll.log; ls -la gpurun_out/prof_ll.ncu-rep
