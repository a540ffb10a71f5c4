"""ctypes access to oracle/_ref/libref_shim.so (the UNMODIFIED reference library behind a batch C shim).

Test infrastructure only: imported by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
--impl reference legs. Never imported by the product package.
"""
import ctypes
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
REF_DIR = os.path.join(os.path.dirname(_HERE), "oracle", "_ref")
_LIB = None


def available():
    return os.path.exists(os.path.join(REF_DIR, "libref_shim.so"))


def lib():
    global _LIB
    if _LIB is None:
        ctypes.CDLL(os.path.join(REF_DIR, "libPolarCode.so"), mode=ctypes.RTLD_GLOBAL)
        L = ctypes.CDLL(os.path.join(REF_DIR, "libref_shim.so"))
        L.ref_frozen_bits.restype = ctypes.c_int
        L.ref_frozen_bits.argtypes = [ctypes.c_int, ctypes.c_int, ctypes.c_float, ctypes.c_char_p, ctypes.c_void_p]
        L.ref_decode_batch2.restype = ctypes.c_double
        L.ref_decode_batch2.argtypes = [ctypes.c_int, ctypes.c_int, ctypes.c_void_p, ctypes.c_int, ctypes.c_char_p, ctypes.c_int,
                                        ctypes.c_char_p, ctypes.c_int, ctypes.c_int, ctypes.c_void_p, ctypes.c_long,
                                        ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_int]
        L.ref_decode_batch.restype = ctypes.c_double
        L.ref_decode_batch.argtypes = [ctypes.c_int, ctypes.c_int, ctypes.c_void_p, ctypes.c_int, ctypes.c_char_p,
                                       ctypes.c_int, ctypes.c_char_p, ctypes.c_int, ctypes.c_int, ctypes.c_void_p,
                                       ctypes.c_long, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_int]
        L.ref_encode_batch.restype = ctypes.c_double
        L.ref_encode_batch.argtypes = [ctypes.c_int, ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_char_p,
                                       ctypes.c_int, ctypes.c_void_p, ctypes.c_long, ctypes.c_void_p, ctypes.c_void_p,
                                       ctypes.c_int, ctypes.c_int]
        L.ref_crc_batch.restype = ctypes.c_int
        L.ref_crc_batch.argtypes = [ctypes.c_int, ctypes.c_char_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_int,
                                    ctypes.c_long, ctypes.c_long, ctypes.c_void_p, ctypes.c_void_p]
        L.ref_decode_soft.restype = ctypes.c_int
        L.ref_decode_soft.argtypes = [ctypes.c_int, ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_void_p, ctypes.c_long,
                                      ctypes.c_void_p, ctypes.c_void_p]
        L.ref_decode_soft_f32.restype = ctypes.c_int
        L.ref_decode_soft_f32.argtypes = [ctypes.c_int, ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_void_p, ctypes.c_long,
                                      ctypes.c_void_p, ctypes.c_void_p]
        L.ref_quantize.restype = ctypes.c_int
        L.ref_quantize.argtypes = [ctypes.c_int, ctypes.c_void_p, ctypes.c_long, ctypes.c_void_p]
        L.ref_puncturer.restype = ctypes.c_long
        L.ref_puncturer.argtypes = [ctypes.c_long, ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_void_p, ctypes.c_long,
                                    ctypes.c_void_p]
        _LIB = L
    return _LIB


def _p(a):
    return a.ctypes.data_as(ctypes.c_void_p)


def frozen_bits(N, K, dsnr, ctype="BB"):
    out = np.zeros(N, dtype=np.uint32)
    n = lib().ref_frozen_bits(N, K, float(dsnr), ctype.encode(), _p(out))
    if n < 0:
        raise RuntimeError("reference frozen_bits failed")
    return out[:n].copy()


def decode_batch(N, L, frozen, kind, llr, crc_bits=-1, crc_type="crc", systematic=True, threads=1, repeat=1,
                 return_time=False, fresh=False):
    """fresh=True: a new decoder object per frame (see oracle/ref_shim.cpp: SclAvxFloat keeps its best path metric
    across decode() calls, which makes float results depend on the decoder's history)"""
    frozen = np.ascontiguousarray(frozen, dtype=np.uint32)
    llr = np.ascontiguousarray(llr)
    is_float = llr.dtype == np.float32
    assert is_float or llr.dtype == np.int8
    B = llr.shape[0]
    K = N - len(frozen)
    kb = (K + 7) // 8
    out = np.zeros((B, kb), dtype=np.uint8)
    ok = np.zeros(B, dtype=np.uint8)
    t = lib().ref_decode_batch2(N, L, _p(frozen), len(frozen), kind.encode(), crc_bits, crc_type.encode(),
                                int(systematic), int(is_float), _p(llr), B, _p(out), _p(ok), threads, repeat, int(fresh))
    if t < 0:
        raise RuntimeError("reference decode failed")
    return (out, ok, t) if return_time else (out, ok)


def decode_soft(N, frozen, llr, systematic=True):
    """Decoder::getSoftCodeword / getSoftInformation of the reference's 'char' Fast-SSC decoder after decode_vector"""
    frozen = np.ascontiguousarray(frozen, dtype=np.uint32)
    llr = np.ascontiguousarray(llr, dtype=np.int8)
    B, K = llr.shape[0], N - len(frozen)
    sc, si = np.zeros((B, N), np.int8), np.zeros((B, K), np.int8)
    if lib().ref_decode_soft(N, _p(frozen), len(frozen), int(systematic), _p(llr), B, _p(sc), _p(si)) < 0:
        raise RuntimeError("reference soft decode failed")
    return sc, si


def decode_soft_f32(N, frozen, llr, systematic=True):
    """the same getters of the 'float' Fast-SSC decoder: float32 [B, N] / [B, K], to be compared as uint32 patterns"""
    frozen = np.ascontiguousarray(frozen, dtype=np.uint32)
    llr = np.ascontiguousarray(llr, dtype=np.float32)
    B, K = llr.shape[0], N - len(frozen)
    sc, si = np.zeros((B, N), np.float32), np.zeros((B, K), np.float32)
    if lib().ref_decode_soft_f32(N, _p(frozen), len(frozen), int(systematic), _p(llr), B, _p(sc), _p(si)) < 0:
        raise RuntimeError("reference float soft decode failed")
    return sc, si


def encode_batch(N, frozen, info, crc_bits=-1, crc_type="crc", systematic=True, threads=1, repeat=1,
                 return_time=False):
    frozen = np.ascontiguousarray(frozen, dtype=np.uint32)
    info = np.ascontiguousarray(info, dtype=np.uint8)
    B = info.shape[0]
    code = np.zeros((B, N // 8), dtype=np.uint8)
    info_out = np.zeros_like(info)
    t = lib().ref_encode_batch(N, _p(frozen), len(frozen), crc_bits, crc_type.encode(), int(systematic), _p(info), B,
                               _p(code), _p(info_out), threads, repeat)
    if t < 0:
        raise RuntimeError("reference encode failed")
    return (code, info_out, t) if return_time else (code, info_out)


def crc_generate(crc_bits, crc_type, data):
    data = np.ascontiguousarray(data, dtype=np.uint8).copy()
    B, nb = data.shape
    if lib().ref_crc_batch(crc_bits, crc_type.encode(), 0, _p(data), nb, 0, B, None, None) < 0:
        raise RuntimeError("reference crc failed")
    return data


def crc_check(crc_bits, crc_type, data):
    data = np.ascontiguousarray(data, dtype=np.uint8).copy()
    B, nb = data.shape
    ok = np.zeros(B, dtype=np.uint8)
    if lib().ref_crc_batch(crc_bits, crc_type.encode(), 1, _p(data), nb, 0, B, _p(ok), None) < 0:
        raise RuntimeError("reference crc failed")
    return ok


def crc_calculate(crc_bits, crc_type, data, bits):
    data = np.ascontiguousarray(data, dtype=np.uint8).copy()
    B, nb = data.shape
    val = np.zeros(B, dtype=np.uint64)
    if lib().ref_crc_batch(crc_bits, crc_type.encode(), 2, _p(data), nb, bits, B, None, _p(val)) < 0:
        raise RuntimeError("reference crc failed")
    return val


def quantize(N, llr_f32):
    x = np.ascontiguousarray(llr_f32, dtype=np.float32)
    B = x.shape[0]
    out = np.zeros((B, N), dtype=np.int8)
    if lib().ref_quantize(N, _p(x), B, _p(out)) < 0:
        raise RuntimeError("reference quantize failed")
    return out


def puncturer(blockLength, frozen, mode, data=None):
    """PolarCode::Puncturer over a batch; mode: 'positions' | 'packed' | 'puncture' | 'depuncture' (dtype of `data`
    picks the template instance). Raises IndexError where the reference throws std::out_of_range."""
    frozen = np.ascontiguousarray(frozen, dtype=np.uint32)
    Np = 1
    while Np < blockLength:
        Np *= 2
    if mode == "positions":
        out = np.zeros(blockLength, dtype=np.uint32)
        r = lib().ref_puncturer(blockLength, _p(frozen), len(frozen), 0, None, 0, _p(out))
    else:
        data = np.ascontiguousarray(data)
        B = data.shape[0]
        t = {np.dtype(np.uint8): 0, np.dtype(np.float32): 1, np.dtype(np.float64): 2}[data.dtype]
        if mode == "packed":
            out, m = np.zeros((B, blockLength // 8), np.uint8), 1
        elif mode == "puncture":
            out, m = np.zeros((B, blockLength), data.dtype), 2 + t
        else:
            out, m = np.zeros((B, Np), data.dtype), 5 + t
        r = lib().ref_puncturer(blockLength, _p(frozen), len(frozen), m, _p(data), B, _p(out))
    if r == -2:
        raise IndexError("Number of required puncturing positions exceeds frozen bit positions!")
    if r < 0:
        raise RuntimeError("reference puncturer failed")
    return out
