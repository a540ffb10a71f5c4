set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_scflip.py -q -x > gpurun_out/r02_c8_scflip.log 2>&1; echo "scflip rc=$?" >> gpurun_out/r02_c8_scflip.log
timeout 600 python -m pytest tests -m gpu -q --deselect tests/test_gpu_scflip.py > gpurun_out/r02_c8_tests.log 2>&1; echo "tests rc=$?" >> gpurun_out/r02_c8_tests.log
{
python - <<'PY'
import sys, ctypes
sys.path.insert(0, '.')
import torch, polar_b200_loader
pcb = polar_b200_loader.load()
N, K = 1024, 512
fz = pcb.frozen_bits(N, K, 0.0)
L_, ptr, check = pcb.capi.lib(), pcb.capi.ptr, pcb.capi.check
for T, B, eb in ((1, 1 << 17, 2.0), (8, 1 << 17, 2.0), (8, 1 << 17, 3.0), (32, 1 << 16, 2.0)):
    dec = pcb.PolarDecoder(N, T, fz, pcb.capi.PC_DEC_SCFLIP_F32)
    llr = torch.empty((B, N), dtype=torch.float32, device="cuda"); info = torch.empty((B, K // 8), dtype=torch.uint8, device="cuda")
    check(L_.pc_awgn_generate(dec._h, 1, 0, eb, 4.0 * 10 ** (eb / 10), 1, ptr(llr), ptr(info), B, None))
    out, ok = dec.decode_batch(llr, return_ok=True); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(3): out, ok = dec.decode_batch(llr, return_ok=True)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 3
    bler = float((out != info).any(dim=1).float().mean())
    print(f"SC-Flip N={N} K={K} T={T} B={B} Eb/N0={eb}: {ms:.3f} ms, {B*K/ms/1e6:.2f} Gbit/s info, BLER {bler:.4f}, detector-false {1-float(ok.float().mean()):.4f}")
PY
python tools/pcsim.py depthfirst -n 256 --l-min 2 --l-max 8 -w 20000000 -o gpurun_out/r02_c8_pcsim > gpurun_out/r02_c8_pcsim.log 2>&1; echo "pcsim rc=$?"
} > gpurun_out/r02_c8_perf.log 2>&1
