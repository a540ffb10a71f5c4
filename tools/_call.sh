set -x
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/r02_c1_smi.txt
python -m pytest tests -m gpu -x -q > gpurun_out/r02_c1_tests.log 2>&1; echo "tests rc=$?" >> gpurun_out/r02_c1_tests.log
python bench.py --steps 20 --warmup 3 > gpurun_out/r02_c1_bench.json 2> gpurun_out/r02_c1_bench.err; echo "bench rc=$?"
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r02_c1_bench_ref.json 2> gpurun_out/r02_c1_bench_ref.err
P="python tools/perf_decoder.py"
NCU="ncu --set full --clock-control none --import-source on"
$P 1024 512 8 char 16 65536 2.0 > gpurun_out/r02_c1_p_scl_i8.log 2>&1 && $NCU -k regex:k_scl_i8 -s 2 -c 1 -f -o gpurun_out/r02_scl_i8_v0 $P 1024 512 8 char 16 65536 2.0 > gpurun_out/ncu1.log 2>&1
$P 1024 512 8 float 16 32768 2.0 > gpurun_out/r02_c1_p_scl_f32.log 2>&1 && $NCU -k regex:k_scl_f32 -s 2 -c 1 -f -o gpurun_out/r02_scl_f32_v0 $P 1024 512 8 float 16 32768 2.0 > gpurun_out/ncu2.log 2>&1
python tools/perf_encoder.py 16384 8192 65536 > gpurun_out/r02_c1_p_enc_c5.log 2>&1 && $NCU -k regex:"k_encode|k_awgn|k_random_info|k_crc_rows" -s 6 -c 8 -f -o gpurun_out/r02_enc_c5_v0 python tools/perf_encoder.py 16384 8192 65536 > gpurun_out/ncu3.log 2>&1
python tools/perf_encoder.py 1024 512 1048576 > gpurun_out/r02_c1_p_enc_c2.log 2>&1 && $NCU -k regex:"k_encode|k_awgn|k_random_info|k_crc_rows" -s 6 -c 8 -f -o gpurun_out/r02_enc_c2_v0 python tools/perf_encoder.py 1024 512 1048576 > gpurun_out/ncu4.log 2>&1
$P 1024 512 1 float 8 262144 2.0 > gpurun_out/r02_c1_p_f32.log 2>&1
ls -la gpurun_out
