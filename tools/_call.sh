set -x
mkdir -p gpurun_out
python -m pytest tests -m gpu -q > gpurun_out/r02_c5_tests.log 2>&1; echo "tests rc=$?" >> gpurun_out/r02_c5_tests.log
python bench.py --steps 20 --warmup 3 > gpurun_out/r02_c5_bench.json 2> gpurun_out/r02_c5_bench.err; echo "bench rc=$?"
python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/r02_c5_bench_ref.json 2> gpurun_out/r02_c5_bench_ref.err
P="python tools/perf_decoder.py"
{
$P 1024 512 1 char 8 1048576 2.0
python - <<'PY'
import sys, time
sys.path.insert(0, '.')
import numpy as np, torch, polar_b200_loader
pcb = polar_b200_loader.load()
N, K, B = 1024, 512, 1 << 18
fz = pcb.frozen_bits(N, K, 0.0)
dec = pcb.PolarDecoder(N, 1, fz, "char")
llr = torch.randint(-60, 100, (B, N), dtype=torch.int8, device="cuda")
for _ in range(2): sc, si = dec.decode_soft_batch(llr)
torch.cuda.synchronize(); e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(5): sc, si = dec.decode_soft_batch(llr)
e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 5
print(f"soft outputs N={N} K={K} B={B}: {ms:.3f} ms, {B*K/ms/1e6:.2f} Gbit/s info")
PY
} > gpurun_out/r02_c5_perf.log 2>&1
ls gpurun_out | head -3
