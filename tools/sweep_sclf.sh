#!/bin/bash
# usage: tools/sweep_sclf.sh "<threads> <ctas>" ...   (N=1024 K=512 L=8 CRC-16, fp32 SCL)
for cfg in "$@"; do
  set -- $cfg
  r=$(PC_SCLF_THREADS=$1 PC_SCLF_CTAS=$2 python tools/perf_decoder.py 1024 512 8 float 16 32768 2.0 2>&1 | tail -1)
  echo "T=$1 ctas=$2 -> $r"
done
