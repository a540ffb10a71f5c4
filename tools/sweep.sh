#!/bin/bash
# usage: tools/sweep.sh "<threads> <ctas> <notmem>" ...   runs bench.py per config (frames = a whole number of
# rounds of the resident threads, so configs are compared without the tail effect) and prints value + ms
for cfg in "$@"; do
  set -- $cfg
  frames=$(python -c "r=148*$1*$2; print(r*max(1,round((1<<20)/r)))")
  r=$(PC_FASTSSC_THREADS=$1 PC_FASTSSC_CTAS=$2 PC_FASTSSC_NOTMEM=$3 python bench.py --steps 10 --warmup 3 --frames $frames --e2e-frames 1024 --no-cpu-baseline --no-secondary 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(round(d['value'],1), round(d['ms_per_step'],3), d['check']['bler'])" 2>&1 | tail -1)
  echo "T=$1 ctas=$2 notmem=$3 frames=$frames -> $r"
done
