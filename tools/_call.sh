set -x
mkdir -p gpurun_out
P="timeout 200 python tools/perf_decoder.py"
{
for k in 1 2 3 4 8 16; do
echo "== lockstep period $k"
PC_SCL_LOCKSTEP=$k $P 1024 512 8 char 16 65536 2.0 | tail -1
PC_SCL_LOCKSTEP=$k $P 1024 512 8 float 16 32768 2.0 | tail -1
done
} > gpurun_out/r02_c21_lockstep.log 2>&1
