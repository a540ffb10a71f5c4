"""ctypes access to oracle/libpc_oracle.so (our scalar C restatement of the reference hot path).

Test infrastructure only (tests/, smoke(), bench cpu_baseline); never imported by the product package.
"""
import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
ORACLE_DIR = os.path.join(os.path.dirname(_HERE), "oracle")
_LIB = None

CRC_KIND = {("crc", 0): 0, ("crc", 6): 6, ("crc", 8): 8, ("crc", 11): 11, ("crc", 16): 16, ("crcnr", 16): 17,
            ("crc", 24): 24, ("crc", 32): 32}


def crc_kind(bits, ctype="crc"):
    nr = "nr" in ctype.lower()
    if bits == 16 and nr:
        return 17
    return bits


def lib():
    global _LIB
    if _LIB is None:
        so = os.path.join(ORACLE_DIR, "libpc_oracle.so")
        srcs = [os.path.join(ORACLE_DIR, f) for f in ("pc_oracle.c", "pc_oracle_scl.c", "pc_oracle.h")]
        if not os.path.exists(so) or os.path.getmtime(so) < max(os.path.getmtime(f) for f in srcs):
            subprocess.check_call(["make", "-C", ORACLE_DIR, "restatement"])
        L = ctypes.CDLL(so)
        vp, ci, cf = ctypes.c_void_p, ctypes.c_int, ctypes.c_float
        L.pco_frozen_bits_bb.restype = ci
        L.pco_frozen_bits_bb.argtypes = [ci, ci, cf, vp]
        L.pco_frozen_bits_be.restype = ci
        L.pco_frozen_bits_be.argtypes = [ci, ci, vp]
        L.pco_f_i8.argtypes = [vp, vp, vp, ci]
        L.pco_g_i8.argtypes = [vp, vp, vp, vp, ci]
        L.pco_quantize.argtypes = [ci, vp, vp]
        L.pco_fastssc_i8_decode.argtypes = [ci, vp, ci, ci, vp, vp, vp]
        L.pco_fastssc_i8_census.argtypes = [ci, vp, ci, vp]
        L.pco_encode.argtypes = [ci, vp, ci, ci, vp, vp]
        L.pco_crc_calculate.restype = ctypes.c_uint64
        L.pco_crc_calculate.argtypes = [ci, vp, ctypes.c_size_t]
        L.pco_crc_generate.argtypes = [ci, vp, ci]
        L.pco_crc_check.restype = ci
        L.pco_crc_check.argtypes = [ci, vp, ci]
        L.pco_scl_i8_decode.restype = ci
        L.pco_scl_i8_decode.argtypes = [ci, vp, ci, ci, ci, ci, vp, vp]
        _LIB = L
    return _LIB


def _p(a):
    return a.ctypes.data_as(ctypes.c_void_p)


def frozen_bits(N, K, dsnr=0.0, ctype="BB"):
    out = np.zeros(N, dtype=np.uint32)
    if "be" in ctype.lower():
        n = lib().pco_frozen_bits_be(N, K, _p(out))
    else:
        n = lib().pco_frozen_bits_bb(N, K, float(dsnr), _p(out))
    return out[:n].copy()


def f_i8(a, b):
    a = np.ascontiguousarray(a, dtype=np.int8); b = np.ascontiguousarray(b, dtype=np.int8)
    o = np.zeros_like(a)
    lib().pco_f_i8(_p(a), _p(b), _p(o), a.size)
    return o


def g_i8(a, b, bits):
    a = np.ascontiguousarray(a, dtype=np.int8); b = np.ascontiguousarray(b, dtype=np.int8)
    bits = np.ascontiguousarray(bits, dtype=np.int8)
    o = np.zeros_like(a)
    lib().pco_g_i8(_p(a), _p(b), _p(bits), _p(o), a.size)
    return o


def quantize(N, x):
    x = np.ascontiguousarray(x, dtype=np.float32).reshape(-1, N)
    o = np.zeros(x.shape, dtype=np.int8)
    for i in range(x.shape[0]):
        lib().pco_quantize(N, _p(x[i]), _p(o[i]))
    return o


def fastssc_i8_decode(N, frozen, llr, systematic=True, soft=False):
    frozen = np.ascontiguousarray(frozen, dtype=np.uint32)
    llr = np.ascontiguousarray(llr, dtype=np.int8).reshape(-1, N)
    K = N - len(frozen)
    out = np.zeros((llr.shape[0], (K + 7) // 8), dtype=np.uint8)
    sb = np.zeros((llr.shape[0], N), dtype=np.int8)
    L = lib()
    for i in range(llr.shape[0]):
        L.pco_fastssc_i8_decode(N, _p(frozen), len(frozen), int(systematic), _p(llr[i]), _p(out[i]), _p(sb[i]))
    return (out, sb) if soft else out


CENSUS_NAMES = ["R0", "R1", "REP", "sREP", "SPC", "sSPC", "DREP", "s01", "s0SPC", "sR1n", "s0R", "sRR", "0SPC", "R1n",
                "0R", "RR"]


def census(N, frozen):
    frozen = np.ascontiguousarray(frozen, dtype=np.uint32)
    c = np.zeros(20, dtype=np.int32)
    lib().pco_fastssc_i8_census(N, _p(frozen), len(frozen), _p(c))
    return dict(zip(CENSUS_NAMES, c.tolist()))


def encode(N, frozen, info, systematic=True):
    frozen = np.ascontiguousarray(frozen, dtype=np.uint32)
    info = np.ascontiguousarray(info, dtype=np.uint8)
    info = info.reshape(-1, info.shape[-1])
    out = np.zeros((info.shape[0], N // 8), dtype=np.uint8)
    for i in range(info.shape[0]):
        lib().pco_encode(N, _p(frozen), len(frozen), int(systematic), _p(info[i]), _p(out[i]))
    return out


def crc_calculate(kind, data, bits):
    data = np.ascontiguousarray(data, dtype=np.uint8)
    return int(lib().pco_crc_calculate(kind, _p(data), bits))


def crc_generate(kind, data):
    data = np.ascontiguousarray(data, dtype=np.uint8).copy()
    data2 = data.reshape(-1, data.shape[-1])
    for i in range(data2.shape[0]):
        lib().pco_crc_generate(kind, _p(data2[i]), data2.shape[1])
    return data


def crc_check(kind, data):
    data = np.ascontiguousarray(data, dtype=np.uint8)
    data2 = data.reshape(-1, data.shape[-1])
    return np.array([lib().pco_crc_check(kind, _p(data2[i]), data2.shape[1]) for i in range(data2.shape[0])],
                    dtype=np.uint8)


def scl_i8_decode(N, frozen, L, llr, crc=8, systematic=True):
    frozen = np.ascontiguousarray(frozen, dtype=np.uint32)
    llr = np.ascontiguousarray(llr, dtype=np.int8).reshape(-1, N)
    K = N - len(frozen)
    out = np.zeros((llr.shape[0], (K + 7) // 8), dtype=np.uint8)
    ok = np.zeros(llr.shape[0], dtype=np.uint8)
    lb = lib()
    for i in range(llr.shape[0]):
        ok[i] = lb.pco_scl_i8_decode(N, _p(frozen), len(frozen), L, crc, int(systematic), _p(llr[i]), _p(out[i]))
    return out, ok
