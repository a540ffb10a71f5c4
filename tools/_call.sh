set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x -k "float or golden or at_scale or adaptive or scl or pcsim" > gpurun_out/r02_c16_tests.log 2>&1; echo "tests rc=$?" >> gpurun_out/r02_c16_tests.log
P="timeout 200 python tools/perf_decoder.py"
{
$P 64 40 4 float 8 262144 2.0
$P 64 40 4 mixed 8 1048576 2.0
$P 256 128 8 float 16 65536 2.0
$P 128 64 4 float 8 262144 2.0
$P 256 128 4 float 8 131072 2.0
$P 1024 512 8 float 16 32768 2.0
} > gpurun_out/r02_c16_perf.log 2>&1
