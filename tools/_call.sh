set -x
mkdir -p gpurun_out
python -m pytest tests -m gpu -q > gpurun_out/r02_c3_tests.log 2>&1; echo "tests rc=$?" >> gpurun_out/r02_c3_tests.log
P="python tools/perf_decoder.py"
{
echo "== Fast-SSC i8 (hints reverted)"
$P 1024 512 1 char 8 1048576 2.0
echo "== SCL f32 shapes"
PC_DEBUG_OCC=1 $P 1024 512 8 float 16 32768 2.0
for t in 128 192 224 320 384; do PC_SCLF_T=$t $P 1024 512 8 float 16 32768 2.0; done
PC_SCLF_VIRT=0 $P 1024 512 8 float 16 32768 2.0
PC_SCL_LOCKSTEP=0 $P 1024 512 8 float 16 32768 2.0
PC_SCL_NOTMEM=1 $P 1024 512 8 float 16 32768 2.0
PC_SCL_NOTMEM=1 PC_SCL_LOCKSTEP=1 $P 1024 512 8 float 16 32768 2.0
$P 512 128 8 float 24 65536 2.0
PC_SCL_NOTMEM=1 $P 512 128 8 float 24 65536 2.0
$P 1024 512 32 float 16 8192 2.0
PC_SCL_NOTMEM=1 $P 1024 512 32 float 16 8192 2.0
$P 1024 512 8 mixed 16 262144 2.0
$P 1024 512 8 achar 16 262144 2.0
$P 1024 512 8 amixed 16 262144 2.0
$P 64 40 4 mixed 8 1048576 2.0
echo "== encoder"
python tools/perf_encoder.py 1024 512 1048576
python tools/perf_encoder.py 16384 8192 65536
python tools/perf_encoder.py 4096 2048 262144
} > gpurun_out/r02_c3_perf.log 2>&1
NCU="ncu --set full --clock-control none --import-source on"
$P 1024 512 8 float 16 32768 2.0 > /dev/null 2>&1 && $NCU -k regex:k_scl_f32 -s 2 -c 1 -f -o gpurun_out/r02_scl_f32_v1 $P 1024 512 8 float 16 32768 2.0 > gpurun_out/ncu1.log 2>&1
python tools/perf_encoder.py 16384 8192 65536 > /dev/null 2>&1 && $NCU -k regex:"k_encode2" -s 4 -c 1 -f -o gpurun_out/r02_enc_c5_v1 python tools/perf_encoder.py 16384 8192 65536 > gpurun_out/ncu3.log 2>&1
$P 1024 512 1 char 8 1048576 2.0 > /dev/null 2>&1 && $NCU -k regex:k_fastssc_i8 -s 2 -c 1 -f -o gpurun_out/r02_fastssc_i8_v5 $P 1024 512 1 char 8 1048576 2.0 > gpurun_out/ncu4.log 2>&1
ls -la gpurun_out | head -40
