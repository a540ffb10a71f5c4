set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -k "scl or at_scale or golden or adaptive or pcsim or fiveg" > gpurun_out/r02_c9_tests.log 2>&1; echo "tests rc=$?" >> gpurun_out/r02_c9_tests.log
P="python tools/perf_decoder.py"
{
$P 1024 512 8 char 16 65536 2.0
PC_SCL_SORT1=1 $P 1024 512 8 char 16 65536 2.0
$P 1024 512 4 char 16 65536 2.0
$P 1024 512 2 char 16 65536 2.0
$P 1024 512 32 char 16 16384 2.0
$P 512 128 8 char 24 65536 2.0
$P 4096 2048 8 char 16 16384 2.0
$P 1024 512 8 achar 16 262144 2.0
} > gpurun_out/r02_c9_perf.log 2>&1
NCU="ncu --set full --clock-control none --import-source on"
$P 1024 512 8 char 16 65536 2.0 > /dev/null 2>&1 && $NCU -k regex:k_scl_i8 -s 2 -c 1 -f -o gpurun_out/r02_scl_i8_v3 $P 1024 512 8 char 16 65536 2.0 > gpurun_out/ncu2.log 2>&1
