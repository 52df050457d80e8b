#!/bin/bash
# tools/gpu_sanitize.sh <memcheck|racecheck|synccheck|initcheck> [cases...]   - ONE tool per gpurun call
mkdir -p gpurun_out
tool=$1; shift
python tools/sanitize_case.py "$@" > gpurun_out/san_plain_$tool.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/san_plain_$tool.log; exit 1; }
tail -4 gpurun_out/san_plain_$tool.log
timeout 1500 compute-sanitizer --tool $tool --print-limit 20 python tools/sanitize_case.py "$@" > gpurun_out/san_$tool.log 2>&1
echo "sanitizer exit $?"; grep -E "ERROR SUMMARY|RACECHECK SUMMARY|generator:|speaker encoder:|vocoder:|done|=========.*(Error|hazard|Invalid|Race)" gpurun_out/san_$tool.log | head -20; tail -3 gpurun_out/san_$tool.log
