set -x
mkdir -p gpurun_out
P="timeout 200 python tools/perf_decoder.py"
{
for i in 1 2; do
echo "== base"; (cd _bisect/base && $P 1024 512 8 float 16 32768 2.0 | tail -1)
echo "== rowhint"; $P 1024 512 8 float 16 32768 2.0 | tail -1
done
echo "== base C3"; (cd _bisect/base && $P 512 128 8 float 24 32768 2.0 | tail -1)
echo "== rowhint C3"; $P 512 128 8 float 24 32768 2.0 | tail -1
echo "== base L32"; (cd _bisect/base && $P 1024 512 32 float 16 8192 2.0 | tail -1)
echo "== rowhint L32"; $P 1024 512 32 float 16 8192 2.0 | tail -1
echo "== base amixed"; (cd _bisect/base && $P 1024 512 8 amixed 16 262144 2.0 | tail -1)
echo "== rowhint amixed"; $P 1024 512 8 amixed 16 262144 2.0 | tail -1
} > gpurun_out/r02_c19_rowhint.log 2>&1
timeout 600 python -m pytest tests -m gpu -q -x -k "scl or adaptive or golden_list or at_scale" > gpurun_out/r02_c19_tests.log 2>&1; echo "tests rc=$?" >> gpurun_out/r02_c19_tests.log
NCU="ncu --set full --clock-control none --import-source on"
$P 1024 512 8 float 16 32768 2.0 > /dev/null 2>&1 && $NCU -k regex:k_scl_f32 -s 2 -c 1 -f -o gpurun_out/r02_scl_f32_v6 $P 1024 512 8 float 16 32768 2.0 > gpurun_out/ncu10.log 2>&1
