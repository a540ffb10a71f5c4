set -x
mkdir -p gpurun_out
python -m pytest tests -m gpu -q -k "scl or at_scale or golden or adaptive or pcsim" > gpurun_out/r02_c4_tests.log 2>&1; echo "tests rc=$?" >> gpurun_out/r02_c4_tests.log
P="python tools/perf_decoder.py"
{
echo "== SCL i8"
$P 1024 512 8 char 16 65536 2.0
$P 1024 512 32 char 16 16384 2.0
$P 512 128 8 char 24 65536 2.0
echo "== SCL f32 shapes"
for t in 256 320 384 512; do PC_SCLF_T=$t PC_DEBUG_OCC=1 $P 1024 512 8 float 16 32768 2.0 2>&1 | tail -2; done
PC_SCLF_T=384 $P 512 128 8 float 24 65536 2.0
PC_SCLF_T=384 $P 1024 512 32 float 16 8192 2.0
} > gpurun_out/r02_c4_perf.log 2>&1
NCU="ncu --set full --clock-control none --import-source on"
PC_SCLF_T=384 $P 1024 512 8 float 16 32768 2.0 > /dev/null 2>&1 && PC_SCLF_T=384 $NCU -k regex:k_scl_f32 -s 2 -c 1 -f -o gpurun_out/r02_scl_f32_v2 $P 1024 512 8 float 16 32768 2.0 > gpurun_out/ncu1.log 2>&1
$P 1024 512 8 char 16 65536 2.0 > /dev/null 2>&1 && $NCU -k regex:k_scl_i8 -s 2 -c 1 -f -o gpurun_out/r02_scl_i8_v2 $P 1024 512 8 char 16 65536 2.0 > gpurun_out/ncu2.log 2>&1
ls -la gpurun_out | head -5
