set -x
mkdir -p gpurun_out
python bench.py --steps 20 --warmup 3 > gpurun_out/r02_bench_v6.json 2> gpurun_out/bench_err.log
python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/r02_bench_ref_v6.json 2> gpurun_out/bench_ref_err.log
python bench.py --kind scl_f32 --steps 20 --warmup 3 --no-secondary > gpurun_out/r02_bench_sclf32_v1.json 2>> gpurun_out/bench_err.log
python bench.py --kind scl_i8 --steps 20 --warmup 3 --no-secondary > gpurun_out/r02_bench_scli8_v1.json 2>> gpurun_out/bench_err.log
