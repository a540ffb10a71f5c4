#!/bin/bash
# usage: tools/sweep_scl.sh kind "<lanes> <smemtop>" ...   (N=1024 K=512 L=8 CRC-16)
kind=$1; shift
for cfg in "$@"; do
  set -- $cfg
  r=$(PC_SCL_LANES=$1 PC_SCL_SMEMTOP=$2 python tools/perf_decoder.py 1024 512 8 $kind 16 65536 2.0 2>&1 | tail -1)
  echo "lanes=$1 top=$2 -> $r"
done
