#!/bin/bash
# times every BASELINE.json config (and the metric's SCL L=8 case) with device-generated LLRs; usage: tools/perf_all.sh
P="python tools/perf_decoder.py"
$P 1024 512 1 char 8 1048576 2.0
$P 1024 512 1 float 8 262144 2.0
$P 1024 512 8 char 16 65536 2.0
$P 1024 512 8 float 16 65536 2.0
$P 1024 512 8 mixed 16 262144 2.0
$P 1024 512 8 achar 16 262144 2.0
$P 1024 512 8 amixed 16 262144 2.0
$P 1024 512 8 achar 16 1048576 3.0
$P 64 40 4 mixed 8 1048576 2.0
$P 64 40 4 float 8 262144 2.0
$P 512 128 8 float 24 65536 2.0
$P 512 128 8 char 24 65536 2.0
$P 1024 512 32 char 16 16384 2.0
$P 1024 512 32 float 16 8192 2.0
$P 16384 8192 1 char 8 65536 2.0
