set -x
mkdir -p gpurun_out
python -m pytest tests -m gpu -q > gpurun_out/r02_c2_tests.log 2>&1; echo "tests rc=$?" >> gpurun_out/r02_c2_tests.log
P="python tools/perf_decoder.py"
{
echo "== SCL i8 shapes"
$P 1024 512 8 char 16 65536 2.0
PC_SCL_T=256 $P 1024 512 8 char 16 65536 2.0
PC_SCL_T=256 PC_SCL_LOCKSTEP=0 $P 1024 512 8 char 16 65536 2.0
PC_SCL_LOCKSTEP=0 $P 1024 512 8 char 16 65536 2.0
PC_SCL_NOTMEM=1 $P 1024 512 8 char 16 65536 2.0
PC_SCL_NOTMEM=1 PC_SCL_LOCKSTEP=1 $P 1024 512 8 char 16 65536 2.0
$P 1024 512 32 char 16 16384 2.0
PC_SCL_NOTMEM=1 $P 1024 512 32 char 16 16384 2.0
$P 512 128 8 char 24 65536 2.0
echo "== Fast-SSC i8 hints"
for h in 0 1 2 3 4 6; do PC_FASTSSC_HINTS=$h $P 1024 512 1 char 8 1048576 2.0; done
PC_FASTSSC_HINTS=0 $P 16384 8192 1 char 8 65536 2.0
PC_FASTSSC_HINTS=3 $P 16384 8192 1 char 8 65536 2.0
echo "== Fast-SSC f32"
PC_F32_G8=0 $P 1024 512 1 float 8 262144 2.0
PC_F32_G8=1 $P 1024 512 1 float 8 262144 2.0
PC_F32_G8=1 $P 256 128 1 float 8 262144 2.0
PC_F32_G8=0 $P 256 128 1 float 8 262144 2.0
PC_F32_G8=1 $P 64 40 4 mixed 8 1048576 2.0
PC_F32_G8=0 $P 64 40 4 mixed 8 1048576 2.0
echo "== encoder"
python tools/perf_encoder.py 1024 512 1048576
python tools/perf_encoder.py 16384 8192 65536
} > gpurun_out/r02_c2_perf.log 2>&1
NCU="ncu --set full --clock-control none --import-source on"
$P 1024 512 8 char 16 65536 2.0 > /dev/null 2>&1 && $NCU -k regex:k_scl_i8 -s 2 -c 1 -f -o gpurun_out/r02_scl_i8_v1 $P 1024 512 8 char 16 65536 2.0 > gpurun_out/ncu1.log 2>&1
$P 1024 512 1 float 8 262144 2.0 > /dev/null 2>&1 && $NCU -k regex:k_fastssc_f32 -s 2 -c 1 -f -o gpurun_out/r02_f32_g8_v1 $P 1024 512 1 float 8 262144 2.0 > gpurun_out/ncu2.log 2>&1
python tools/perf_encoder.py 1024 512 1048576 > /dev/null 2>&1 && $NCU -k regex:"k_encode2" -s 4 -c 1 -f -o gpurun_out/r02_enc_c2_v1 python tools/perf_encoder.py 1024 512 1048576 > gpurun_out/ncu3.log 2>&1
PC_FASTSSC_HINTS=0 $P 1024 512 1 char 8 1048576 2.0 > /dev/null 2>&1 && PC_FASTSSC_HINTS=0 $NCU -k regex:k_fastssc_i8 -s 2 -c 1 -f -o gpurun_out/r02_fastssc_i8_h0 $P 1024 512 1 char 8 1048576 2.0 > gpurun_out/ncu4.log 2>&1
PC_FASTSSC_HINTS=3 $NCU -k regex:k_fastssc_i8 -s 2 -c 1 -f -o gpurun_out/r02_fastssc_i8_h3 $P 1024 512 1 char 8 1048576 2.0 > gpurun_out/ncu5.log 2>&1
ls -la gpurun_out | head -40
