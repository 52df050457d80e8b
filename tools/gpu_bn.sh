#!/bin/bash
# train-mode BatchNorm: GPU tests + timing; then the vocoder tests that -x cut off, and ncu captures of the vocoder loop
mkdir -p gpurun_out
echo "== bn_train tests"; timeout 600 python -m pytest tests/test_gpu_bn_train.py -m gpu -q -s 2>&1 | tail -25
echo "== bn_train bench"; timeout 300 python tools/bench_bn_train.py > gpurun_out/bn_train_bench.json 2> gpurun_out/bn_train_bench.err; echo "exit $?"; cat gpurun_out/bn_train_bench.json; tail -3 gpurun_out/bn_train_bench.err
echo "== remaining gpu tests"; timeout 900 python -m pytest tests/test_gpu_vocoder.py tests/test_gpu_parity.py -m gpu -q 2>&1 | tail -5
echo "== vocoder ncu"
export AVC_VOC_NOCOOP=1
timeout 200 python tools/profile_voc.py > gpurun_out/voc_plain.log 2>&1; echo "plain exit $?"; tail -2 gpurun_out/voc_plain.log
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 40 --csv --log-file gpurun_out/voc_launches.csv python tools/profile_voc.py > gpurun_out/voc_ncu1.log 2>&1; echo "launch list exit $?"
STEPS=60 timeout 900 ncu --set full --clock-control none --import-source on -k regex:voc_generate -c 2 -o gpurun_out/prof_voc -f python tools/profile_voc.py > gpurun_out/voc_ncu2.log 2>&1; echo "full exit $?"; tail -3 gpurun_out/voc_ncu2.log
ls -la gpurun_out
