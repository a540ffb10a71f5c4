#!/bin/bash
# usage: tools/sweep.sh "<split> <threads>" ...   runs bench.py per config and prints value + ms
for cfg in "$@"; do
  set -- $cfg
  r=$(PC_FASTSSC_SPLIT=$1 PC_FASTSSC_THREADS=$2 python bench.py --steps 5 --warmup 3 --frames 524288 --e2e-frames 1024 --no-cpu-baseline 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(round(d['value'],1), round(d['ms_per_step'],3), d['check'])" 2>&1 | tail -1)
  echo "split=$1 T=$2 -> $r"
done
