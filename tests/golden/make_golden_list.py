#!/usr/bin/env python
"""Generates tests/golden/reference_vectors_list.npz from the UNMODIFIED reference (oracle/_ref): the list decoders
(SclFipChar, SclAvxFloat), the float Fast-SSC decoder (FastSscAvxFloat) and the adaptive 'mixed' decoder
(AdaptiveFloat). Run in the build container: python tests/golden/make_golden_list.py

Inputs are noisy transmitted codewords (so that the detector accepts some list path). Float list decoding uses a fresh
reference decoder object per frame: SclAvxFloat carries its best path metric into the next decode() call
(scl_avx_float.cpp:103-109), a history dependence that a batch decoder cannot and should not reproduce (DESIGN.md par. 4).
These fixtures pin oracle/pc_oracle_scl.c on the CPU and every list / float kernel on the GPU box."""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import refshim  # noqa: E402

# (N, K, L, crc bits, crc type, frames)
CHAR_CASES = [(1024, 512, 8, 16, "crc", 64), (1024, 512, 32, 16, "crc", 24), (1024, 512, 2, 8, "crc", 48),
              (256, 128, 4, 8, "crc", 96), (128, 64, 8, 8, "crc", 96), (64, 40, 4, 8, "crc", 128),
              (2048, 1024, 8, 24, "crc", 16), (512, 128, 8, 16, "crcnr", 64), (32, 16, 4, 8, "crc", 128)]
FLOAT_CASES = [(1024, 512, 8, 16, "crc", 32), (256, 128, 4, 8, "crc", 64), (64, 40, 4, 8, "crc", 96),
               (512, 128, 8, 24, "crc", 32), (1024, 512, 1, 8, "crc", 32), (128, 64, 1, 8, "crc", 96),
               (2048, 1800, 1, 8, "crc", 8)]
MIXED_CASES = [(1024, 512, 8, 16, "crc", 32), (256, 128, 4, 8, "crc", 64)]
# AdaptiveChar / AdaptiveMixed (pcsim precision 8 / 832, simulator.cpp:721-733): float LLRs = 10 * y (setup.cpp:58)
PCSIM_CASES = [("adaptive_char", 1024, 512, 8, 16, "crc", 48), ("adaptive_char", 256, 128, 4, 8, "crc", 64),
               ("adaptive_mixed", 1024, 512, 8, 16, "crc", 32), ("adaptive_mixed", 128, 64, 4, 8, "crc", 64)]


def transmit(rng, N, K, fz, crc_bits, crc_type, B, ebn0):
    info = rng.integers(0, 256, (B, (K + 7) // 8), dtype=np.uint8)
    if K % 8:
        info[:, -1] &= (0xFF00 >> (K % 8)) & 0xFF
    code, _ = refshim.encode_batch(N, fz, info, crc_bits=crc_bits, crc_type=crc_type)
    x = 1.0 - 2.0 * np.unpackbits(code, axis=1, bitorder="big")[:, :N].astype(np.float32)
    sigma = 1.0 / np.sqrt(2.0 * (K / N) * 10.0 ** (ebn0 / 10.0))
    return (x + sigma * rng.standard_normal((B, N))).astype(np.float32)


def main():
    rng = np.random.default_rng(20261105)
    out = {}
    for N, K, L, cb, ct, B in CHAR_CASES:
        fz = refshim.frozen_bits(N, K, 0.0)
        y = np.concatenate([transmit(rng, N, K, fz, cb, ct, B // 2, 1.0), transmit(rng, N, K, fz, cb, ct, B - B // 2, 2.5)])
        llr = np.clip(np.rint(10.0 * y), -128, 127).astype(np.int8)
        info, ok = refshim.decode_batch(N, L, fz, "char", llr, crc_bits=cb, crc_type=ct)
        key = f"sclc_{N}_{K}_{L}_{cb}_{ct}"
        out[key + "_frozen"], out[key + "_llr"], out[key + "_info"], out[key + "_ok"] = fz, llr, info, ok
    for kind, cases in (("float", FLOAT_CASES), ("mixed", MIXED_CASES)):
        for N, K, L, cb, ct, B in cases:
            fz = refshim.frozen_bits(N, K, 0.0)
            y = np.concatenate([transmit(rng, N, K, fz, cb, ct, B // 2, 1.0), transmit(rng, N, K, fz, cb, ct, B - B // 2, 2.5)])
            llr = (np.float32(4.0) * y).astype(np.float32)
            llr[B // 4: B // 2] = np.rint(llr[B // 4: B // 2])      # integer valued rows: exact metric ties
            info, ok = refshim.decode_batch(N, L, fz, kind, llr, crc_bits=cb, crc_type=ct, fresh=True)
            key = f"{kind}_{N}_{K}_{L}_{cb}_{ct}"
            out[key + "_frozen"], out[key + "_llr"], out[key + "_info"], out[key + "_ok"] = fz, llr.astype(np.float32), info, ok
    for kind, N, K, L, cb, ct, B in PCSIM_CASES:
        fz = refshim.frozen_bits(N, K, 0.0)
        y = np.concatenate([transmit(rng, N, K, fz, cb, ct, B // 2, 1.0), transmit(rng, N, K, fz, cb, ct, B - B // 2, 2.5)])
        llr = (np.float32(10.0) * y).astype(np.float32)
        info, ok = refshim.decode_batch(N, L, fz, kind, llr, crc_bits=cb, crc_type=ct, fresh=True)
        key = f"{kind.replace('_', '')}_{N}_{K}_{L}_{cb}_{ct}"
        out[key + "_frozen"], out[key + "_llr"], out[key + "_info"], out[key + "_ok"] = fz, llr, info, ok
    np.savez_compressed(os.path.join(HERE, "reference_vectors_list.npz"), **out)
    print("wrote", len(out), "arrays")


if __name__ == "__main__":
    main()
