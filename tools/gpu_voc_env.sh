#!/bin/bash
# usage: gpu_voc_env.sh "ENV1=a ENV2=b" "ENV1=c" ...   timings of the vocoder loop under each environment
for e in "$@"; do echo "== $e"; env $e timeout 300 python tools/voc_probe.py 2>&1 | tail -4; done
