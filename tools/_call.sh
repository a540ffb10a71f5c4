set -x
mkdir -p gpurun_out
python -m pytest tests -m gpu -q > gpurun_out/r02_c6_tests.log 2>&1; echo "tests rc=$?" >> gpurun_out/r02_c6_tests.log
P="python tools/perf_decoder.py"
{
echo "== SCL"
$P 1024 512 8 char 16 65536 2.0
$P 1024 512 8 float 16 32768 2.0
$P 1024 512 32 char 16 16384 2.0
$P 1024 512 32 float 16 8192 2.0
$P 512 128 8 char 24 65536 2.0
$P 512 128 8 float 24 65536 2.0
$P 1024 512 8 achar 16 262144 2.0
$P 1024 512 8 achar 16 1048576 3.0
$P 1024 512 8 mixed 16 262144 2.0
$P 64 40 4 mixed 8 1048576 2.0
$P 64 40 4 float 8 262144 2.0
$P 16384 8192 1 char 8 65536 2.0
echo "== encoder / generator"
python tools/perf_encoder.py 1024 512 1048576
python tools/perf_encoder.py 16384 8192 65536
} > gpurun_out/r02_c6_perf.log 2>&1
ls gpurun_out | head -3
