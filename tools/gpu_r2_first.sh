#!/bin/bash
# round 2, first GPU pass: all gpu tests, smoke, smoke under ncu (does the cooperative launch fall back?), bench
mkdir -p gpurun_out
echo "== all gpu tests"; timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/t_all.log 2>&1; echo "exit $?"; tail -15 gpurun_out/t_all.log
echo "== smoke"; timeout 300 python __graft_entry__.py smoke 2>&1 | tail -4
echo "== smoke under ncu (driver-style launch list)"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 1000 --csv --log-file gpurun_out/smoke_launches.csv python __graft_entry__.py smoke > gpurun_out/smoke_ncu.log 2>&1; echo "ncu exit $?"; tail -5 gpurun_out/smoke_ncu.log
grep -c "ll_lstm_kernel\|voc_generate" gpurun_out/smoke_launches.csv; grep "ll_lstm_kernel\|voc_generate" gpurun_out/smoke_launches.csv | head -5 | cut -c1-300
echo "== bench"; timeout 900 python bench.py > gpurun_out/b_r2_first.log 2> gpurun_out/b_r2_first.err; echo "exit $?"; tail -3 gpurun_out/b_r2_first.err; cat gpurun_out/b_r2_first.log | cut -c1-6000
echo "== reference arm"; timeout 900 python bench.py --impl reference --steps 3 --warmup 1 | cut -c1-1200
