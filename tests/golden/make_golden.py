#!/usr/bin/env python
"""Generates tests/golden/*.npz from the UNMODIFIED reference (oracle/_ref, built from /root/reference by
oracle/Makefile). Run in the build container: python tests/golden/make_golden.py
Fixtures: seeded inputs + the reference's outputs for the hot path (construction, Fast-SSC char decode, encoder,
detectors, quantiser). They pin oracle/pc_oracle.c (CPU tests) and the CUDA path (GPU tests) on boxes where the
reference sources are absent."""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import refshim  # noqa: E402


def llr_i8(rng, B, N, ebn0, rate, amp):
    sigma = 1.0 / np.sqrt(2.0 * rate * 10.0 ** (ebn0 / 10.0))
    y = 1.0 + sigma * rng.standard_normal((B, N), dtype=np.float32)
    return np.clip(np.rint(amp * y), -128, 127).astype(np.int8)


def main():
    rng = np.random.default_rng(20261104)
    out = {}
    # construction
    for N, K, d, t in [(8, 4, 0.0, "BB"), (128, 64, 0.0, "BB"), (1024, 512, 0.0, "BB"), (64, 40, 1.0, "BB"),
                       (512, 128, 0.0, "BB"), (16384, 8192, 0.0, "BB"), (256, 128, 0.0, "BE"), (64, 32, 0.0, "BE")]:
        out[f"frozen_{t}_{N}_{K}_{d}"] = refshim.frozen_bits(N, K, d, t)
    # Fast-SSC char decode (systematic, CRC-8 default detector)
    for N, K, B in [(1024, 512, 96), (64, 40, 200), (8, 4, 64), (256, 64, 128), (2048, 1800, 24), (16384, 8192, 6)]:
        fz = refshim.frozen_bits(N, K, 0.0)
        llr = np.concatenate([llr_i8(rng, B // 2, N, 1.5, K / N, 10.0), llr_i8(rng, B // 4, N, 3.0, K / N, 40.0),
                              rng.integers(-128, 128, (B - B // 2 - B // 4, N), dtype=np.int8)])
        info, ok = refshim.decode_batch(N, 1, fz, "char", llr)
        out[f"dec_{N}_{K}_frozen"] = fz
        out[f"dec_{N}_{K}_llr"] = llr
        out[f"dec_{N}_{K}_info"] = info
        out[f"dec_{N}_{K}_ok"] = ok
    # non-systematic (reference is only self-consistent for N >= 256)
    fz = refshim.frozen_bits(1024, 512, 0.0)
    llr = llr_i8(rng, 48, 1024, 2.0, 0.5, 10.0)
    info, ok = refshim.decode_batch(1024, 1, fz, "char", llr, systematic=False)
    out["decns_1024_512_llr"], out["decns_1024_512_info"], out["decns_1024_512_ok"] = llr, info, ok
    # encoder
    for N, K in [(1024, 512), (256, 128), (4096, 3072), (16384, 8192)]:
        fz = refshim.frozen_bits(N, K, 0.0)
        info = rng.integers(0, 256, (16, K // 8), dtype=np.uint8)
        for syst in (1, 0):
            code, _ = refshim.encode_batch(N, fz, info, crc_bits=0, systematic=bool(syst))
            out[f"enc_{N}_{K}_{syst}_info"], out[f"enc_{N}_{K}_{syst}_code"] = info, code
    # detectors
    data = rng.integers(0, 256, (32, 24), dtype=np.uint8)
    out["crc_data"] = data
    for bits, typ in [(6, "crc"), (8, "crc"), (11, "crc"), (16, "crc"), (16, "crcnr"), (24, "crc"), (32, "crc")]:
        cb = {6: 1, 8: 1, 11: 2, 16: 2, 24: 3, 32: 4}[bits]
        full = np.zeros((32, 24 + cb), np.uint8)
        full[:, :24] = data
        out[f"crc_{bits}_{typ}_gen"] = refshim.crc_generate(bits, typ, full)
        if bits != 32:
            out[f"crc_{bits}_{typ}_calc35"] = refshim.crc_calculate(bits, typ, data, 35)
    # quantiser
    x = (rng.standard_normal((8, 64)) * 70).astype(np.float32)
    x[0, :16] = [0.5, 1.5, 2.5, -0.5, -1.5, -2.5, 127.5, -128.5, 1e10, -1e10, 3e9, np.nan, np.inf, -np.inf, 126.5, -127.5]
    out["quant_in"] = x
    out["quant_out_64"] = refshim.quantize(64, x)
    out["quant_out_16"] = refshim.quantize(16, x.reshape(-1, 16))
    np.savez_compressed(os.path.join(HERE, "reference_vectors.npz"), **out)
    print("wrote", len(out), "arrays")


if __name__ == "__main__":
    main()
