#!/bin/bash
# vocoder: probe (parity numbers + timings) for each cluster size given, then the GPU tests
for c in ${CLUSTERS:-8}; do echo "== AVC_VOC_CLUSTER=$c"; AVC_VOC_CLUSTER=$c timeout 300 python tools/voc_probe.py 2>&1 | tail -${TAILN:-7}; done
timeout 500 python -m pytest tests/test_gpu_vocoder.py -m gpu -q 2>&1 | tail -3
