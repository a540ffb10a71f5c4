set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -q -x -k "device or binding or multi" > gpurun_out/r02_c20_tests.log 2>&1; echo "tests rc=$?" >> gpurun_out/r02_c20_tests.log
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/r02_bench_2gpu_v2.json 2> gpurun_out/bench2_err.log
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29534 bench.py --impl reference --gpus 2 --steps 3 --warmup 1 > gpurun_out/r02_bench_ref_2gpu_v2.json 2>> gpurun_out/bench2_err.log
