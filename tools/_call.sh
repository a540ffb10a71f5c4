set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x -k "soft or float or cxx or host or golden" > gpurun_out/r02_c18_tests.log 2>&1; echo "tests rc=$?" >> gpurun_out/r02_c18_tests.log
python - > gpurun_out/r02_c18_softperf.log 2>&1 <<'PY'
import sys, os, torch, ctypes
sys.path.insert(0, os.getcwd())
import polar_b200_loader
pcb = polar_b200_loader.load()
N, K, B = 1024, 512, 131072
fz = pcb.frozen_bits(N, K, 0.0)
for kind, dt in (("float", torch.float32), ("char", torch.int8)):
    dec = pcb.PolarDecoder(N, 1, fz, kind)
    isf = kind == "float"
    llr = torch.empty((B, N), dtype=dt, device="cuda"); info = torch.empty((B, K // 8), dtype=torch.uint8, device="cuda")
    amp = 4.0 * 10 ** 0.2 if isf else 10.0
    pcb.capi.check(pcb.capi.lib().pc_awgn_generate(dec._h, 1, 0, ctypes.c_float(2.0), ctypes.c_float(amp), int(isf), pcb.capi.ptr(llr), pcb.capi.ptr(info), B, None))
    sc, si = dec.decode_soft_batch(llr); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5): sc, si = dec.decode_soft_batch(llr)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 5
    print(f"soft outputs {kind} N={N} K={K} B={B}: {ms:.3f} ms/launch, {B*K/ms/1e6:.2f} Gbit/s info")
PY
