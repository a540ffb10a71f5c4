#!/usr/bin/env python
"""times one decoder configuration on the GPU with device-generated LLRs: tools/perf_decoder.py N K L kind crc frames ebn0
kind: char | float | mixed | achar | amixed"""
import ctypes, sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import polar_b200_loader
pcb = polar_b200_loader.load()
N, K, L, kind, crc, B, ebn0 = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3]), sys.argv[4], int(sys.argv[5]), int(sys.argv[6]), float(sys.argv[7])
fz = pcb.frozen_bits(N, K, 0.0)
# achar / amixed: AdaptiveChar / AdaptiveMixed (pcsim -p 8 / -p 832 with L > 1), LLRs scaled by 10 like pcsim
gpu_kind = {"achar": pcb.capi.PC_DEC_ADAPTIVE_I8, "amixed": pcb.capi.PC_DEC_ADAPTIVE_MIXED}.get(kind, kind)
dec = pcb.PolarDecoder(N, L, fz, gpu_kind)
dec.setErrorDetection(crc, "crc")
isf = "char" not in kind
llr = torch.empty((B, N), dtype=torch.float32 if isf else torch.int8, device="cuda")
info = torch.empty((B, (K + 7) // 8), dtype=torch.uint8, device="cuda")
Lb = pcb.capi.lib(); ptr = pcb.capi.ptr
amp = 4.0 * 10 ** (ebn0 / 10) if (isf and kind != "amixed") else 10.0
pcb.capi.check(Lb.pc_awgn_generate(dec._h, 1, 0, ebn0, amp, int(isf), ptr(llr), ptr(info), B, None))
for _ in range(3):  # the first launches after a pause run 20-40 % slower (tools/launch_times.py)
    out, ok = dec.decode_batch(llr, return_ok=True)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
reps = 10
e0.record()
for _ in range(reps):
    out, ok = dec.decode_batch(llr, return_ok=True)
e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / reps
counts = torch.zeros(4, dtype=torch.int64, device="cuda")
pcb.capi.check(Lb.pc_count_errors(ptr(info), ptr(out), ptr(ok), K, B, ptr(counts), None)); torch.cuda.synchronize()
c = counts.tolist()
print(f"N={N} K={K} L={L} {kind} crc={crc} B={B} Eb/N0={ebn0}: {ms:.3f} ms/launch, {B*K/ms/1e6:.2f} Gbit/s info, {B/ms/1e3:.2f} Mframes/s, BLER {c[1]/c[0]:.4f}, RER {c[3]/c[0]:.4f}")
