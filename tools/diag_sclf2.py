#!/usr/bin/env python
"""dumps the first frames on which k_scl_f32 and SclAvxFloat disagree (small codes) to gpurun_out/sclf_mismatch.npz"""
import ctypes, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np, torch
import polar_b200_loader, refshim
pcb = polar_b200_loader.load()
L_, ptr, check = pcb.capi.lib(), pcb.capi.ptr, pcb.capi.check
B = 1 << 17
save = {}
for N, K, L in ((32, 16, 8), (32, 16, 2), (64, 32, 2), (16, 8, 4), (32, 16, 4)):
    fz = pcb.frozen_bits(N, K, 0.0)
    dec = pcb.PolarDecoder(N, L, fz, "float"); dec.setErrorDetection(0, "crc")
    llr = torch.empty((B, N), dtype=torch.float32, device="cuda"); info = torch.empty((B, (K + 7) // 8), dtype=torch.uint8, device="cuda")
    check(L_.pc_awgn_generate(dec._h, ctypes.c_uint64(1), ctypes.c_uint64(0), ctypes.c_float(0.0), ctypes.c_float(4.0), 1, ptr(llr), ptr(info), B, None))
    h = llr.cpu().numpy()
    out, ok = dec.decode_batch(h, return_ok=True)
    exp, eok = refshim.decode_batch(N, L, np.array(fz, np.uint32), "float", h, crc_bits=0, crc_type="crc", threads=16)
    bad = np.flatnonzero((out != exp).any(axis=1))
    print(N, K, L, "mismatches", len(bad), flush=True)
    if len(bad):
        k = f"{N}_{K}_{L}"
        save[k + "_llr"] = h[bad[:16]]; save[k + "_gpu"] = out[bad[:16]]; save[k + "_ref"] = exp[bad[:16]]; save[k + "_fz"] = np.array(fz, np.uint32)
np.savez("gpurun_out/sclf_mismatch.npz", **save)
