#!/usr/bin/env python
"""diagnostic: where does the fp32 SCL kernel part from SclAvxFloat? mismatch counts over code sizes / list sizes / LLR kinds"""
import ctypes, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np, torch
import polar_b200_loader, refshim
pcb = polar_b200_loader.load()
L_, ptr, check = pcb.capi.lib(), pcb.capi.ptr, pcb.capi.check
B = 16384
for N, K in ((16, 8), (32, 16), (64, 32), (128, 64), (256, 128), (1024, 512)):
    for L in (2, 8):
        for crc in (0, 8):
            fz = pcb.frozen_bits(N, K, 0.0)
            dec = pcb.PolarDecoder(N, L, fz, "float"); dec.setErrorDetection(crc, "crc")
            llr = torch.empty((B, N), dtype=torch.float32, device="cuda"); info = torch.empty((B, (K + 7) // 8), dtype=torch.uint8, device="cuda")
            check(L_.pc_awgn_generate(dec._h, ctypes.c_uint64(1), ctypes.c_uint64(0), ctypes.c_float(0.0), ctypes.c_float(4.0), 1, ptr(llr), ptr(info), B, None))
            h = llr.cpu().numpy()
            res = []
            for name, x in (("cont", h), ("rint", np.rint(h).astype(np.float32)), ("x8rint", (np.rint(h * 8) / 8).astype(np.float32))):
                out, ok = dec.decode_batch(x, return_ok=True)
                exp, eok = refshim.decode_batch(N, L, np.array(fz, np.uint32), "float", x, crc_bits=crc, crc_type="crc", threads=16)
                bad = (out != exp).any(axis=1)
                res.append(f"{name}:{int(bad.sum())}/{int((ok != eok).sum())}")
            print(f"N={N} K={K} L={L} crc={crc}: " + "  ".join(res), flush=True)
