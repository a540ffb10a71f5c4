#!/bin/bash
# usage: tools/sweep.sh "<threads> <ctas> <notmem>" ...   runs bench.py per config and prints value + ms
for cfg in "$@"; do
  set -- $cfg
  r=$(PC_FASTSSC_THREADS=$1 PC_FASTSSC_CTAS=$2 PC_FASTSSC_NOTMEM=$3 python bench.py --steps 5 --warmup 3 --frames 524288 --e2e-frames 1024 --no-cpu-baseline 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(round(d['value'],1), round(d['ms_per_step'],3), d['check'])" 2>&1 | tail -1)
  echo "T=$1 ctas=$2 notmem=$3 -> $r"
done
