#!/usr/bin/env python
"""times the SC-Flip decoder (DepthFirst) with device-generated LLRs: tools/perf_scflip.py N K trials frames ebn0"""
import ctypes, sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import polar_b200_loader
pcb = polar_b200_loader.load()
N, K, T, B, ebn0 = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4]), float(sys.argv[5])
capi = pcb.capi
fz = pcb.frozen_bits(N, K, 0.0)
dec = pcb.PolarDecoder(N, T, fz, capi.PC_DEC_SCFLIP_F32)
dec.setErrorDetection(8, "crc")
llr = torch.empty((B, N), dtype=torch.float32, device="cuda")
info = torch.empty((B, (K + 7) // 8), dtype=torch.uint8, device="cuda")
capi.check(capi.lib().pc_awgn_generate(dec._h, 1, 0, ctypes.c_float(ebn0), ctypes.c_float(4.0 * 10 ** (ebn0 / 10)), 1, capi.ptr(llr), capi.ptr(info), B, None))
for _ in range(2):
    out, ok = dec.decode_batch(llr, return_ok=True)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
reps = 3
e0.record()
for _ in range(reps):
    out, ok = dec.decode_batch(llr, return_ok=True)
e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / reps
counts = torch.zeros(4, dtype=torch.int64, device="cuda")
capi.check(capi.lib().pc_count_errors(capi.ptr(info), capi.ptr(out), capi.ptr(ok), K, B, capi.ptr(counts), None)); torch.cuda.synchronize()
c = counts.tolist()
print(f"SC-Flip N={N} K={K} T={T} B={B} Eb/N0={ebn0}: {ms:.3f} ms, {B * K / ms / 1e6:.2f} Gbit/s info, BLER {c[1] / c[0]:.4f}, detector-false {c[3] / c[0]:.4f}")
