#!/usr/bin/env python
"""times the packed encoder and the channel generator with device buffers: tools/perf_encoder.py N K frames"""
import ctypes, sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import polar_b200_loader
pcb = polar_b200_loader.load()
N, K, B = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
fz = pcb.frozen_bits(N, K, 0.0)
Lb = pcb.capi.lib(); ptr = pcb.capi.ptr; check = pcb.capi.check
for syst in (True, False):
    for crc in (0, 8):
        enc = pcb.PolarEncoder(N, fz); enc.setSystematic(syst); enc.setErrorDetection(crc, "crc")
        info = torch.randint(0, 256, (B, K // 8), dtype=torch.uint8, device="cuda")
        code = torch.empty((B, N // 8), dtype=torch.uint8, device="cuda")
        for _ in range(3):
            check(Lb.pc_encode_batch(enc._h, ptr(info), ptr(code), None, B, None))
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        reps = 20
        e0.record()
        for _ in range(reps):
            check(Lb.pc_encode_batch(enc._h, ptr(info), ptr(code), None, B, None))
        e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / reps
        byts = B * (K // 8 + N // 8)
        print(f"encode N={N} K={K} B={B} systematic={syst} crc={crc}: {ms:.3f} ms, {B*N/ms/1e6:.1f} Gbit/s coded, {byts/ms/1e6:.0f} GB/s algorithmic ({byts/ms/1e6/6549*100:.1f} % of 6549)")
dec = pcb.PolarDecoder(N, 1, fz, "char")
llr = torch.empty((B, N), dtype=torch.int8, device="cuda"); info = torch.empty((B, K // 8), dtype=torch.uint8, device="cuda")
for isf, buf in ((0, llr), (1, torch.empty((B, N), dtype=torch.float32, device="cuda"))):
    for _ in range(2):
        check(Lb.pc_awgn_generate(dec._h, ctypes.c_uint64(1), ctypes.c_uint64(0), ctypes.c_float(2.0), ctypes.c_float(10.0), isf, ptr(buf), ptr(info), B, None))
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10):
        check(Lb.pc_awgn_generate(dec._h, ctypes.c_uint64(1), ctypes.c_uint64(0), ctypes.c_float(2.0), ctypes.c_float(10.0), isf, ptr(buf), ptr(info), B, None))
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 10
    out_bytes = B * N * (4 if isf else 1)
    print(f"awgn_generate N={N} B={B} float={isf}: {ms:.3f} ms, {B/ms/1e3:.1f} Mframes/s, LLR write {out_bytes/ms/1e6:.0f} GB/s")
