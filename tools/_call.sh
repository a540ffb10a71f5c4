set -x
mkdir -p gpurun_out
P="timeout 300 python tools/perf_decoder.py"
$P 1024 512 8 amixed 16 262144 2.0 > gpurun_out/c14_a.log 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/c14_amixed_launches.csv $P 1024 512 8 amixed 16 262144 2.0 > gpurun_out/c14_b.log 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/c14_mixed_launches.csv $P 1024 512 8 mixed 16 262144 2.0 > gpurun_out/c14_c.log 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none -c 40 --csv --log-file gpurun_out/c14_f64_launches.csv $P 64 40 4 float 8 262144 2.0 > gpurun_out/c14_d.log 2>&1
$P 64 40 4 float 8 262144 2.0 >> gpurun_out/c14_a.log 2>&1
$P 512 256 1 float 8 262144 2.0 >> gpurun_out/c14_a.log 2>&1
PC_F32_SM_TM=1 $P 512 256 1 float 8 262144 2.0 >> gpurun_out/c14_a.log 2>&1
PC_F32_SM=0 $P 512 256 1 float 8 262144 2.0 >> gpurun_out/c14_a.log 2>&1
