"""ctypes binding of libpolar_b200.so (include/polar_b200.h) -- the only way Python reaches the CUDA path.

There is no CPU fallback: if the shared library is missing this module raises on first use, and every
compute entry point fails with PC_ERR_CUDA when no GPU is usable.
"""
import ctypes
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libpolar_b200.so")

PC_DEC_FASTSSC_I8, PC_DEC_FASTSSC_F32, PC_DEC_SCL_I8, PC_DEC_SCL_F32, PC_DEC_ADAPTIVE_F32, PC_DEC_ADAPTIVE_I8, \
    PC_DEC_ADAPTIVE_MIXED, PC_DEC_SCFLIP_F32 = range(8)
PC_OK, PC_ERR_ARG, PC_ERR_CUDA, PC_ERR_UNSUPPORTED = range(4)

FLOAT_LLR_KINDS = (PC_DEC_FASTSSC_F32, PC_DEC_SCL_F32, PC_DEC_ADAPTIVE_F32, PC_DEC_ADAPTIVE_MIXED, PC_DEC_SCFLIP_F32)


def kind_takes_float(kind):
    """True for the decoder kinds pcsim feeds with float LLRs (simulator.cpp:703-765): the float decoders and
    AdaptiveMixed, whose list pass needs the unquantised floats (adaptive_mixed.cpp:36-47)"""
    return int(kind) in FLOAT_LLR_KINDS


def pcsim_amplification(kind, ebn0_db, fixed=10.0):
    """pcsim's LLR scaling (simulator.cpp:370-379, setup.cpp:58): the fixed amplification (default 10) for the int8
    decoders and for AdaptiveMixed (-p 832), 4 * Eb/N0 (linear) for the float decoders"""
    if kind_takes_float(kind) and int(kind) != PC_DEC_ADAPTIVE_MIXED:
        return 4.0 * 10.0 ** (float(ebn0_db) / 10.0)
    return float(fixed)


_vp, _i, _sz, _u64 = ctypes.c_void_p, ctypes.c_int, ctypes.c_size_t, ctypes.c_uint64

# name -> (restype, argtypes); mirrors include/polar_b200.h one to one (tests check the export list)
SIGNATURES = {
    "pc_puncturer_create": (_vp, [ctypes.c_size_t, _vp, _i]),
    "pc_puncturer_destroy": (None, [_vp]),
    "pc_puncturer_block_length": (ctypes.c_size_t, [_vp]),
    "pc_puncturer_parent_block_length": (ctypes.c_size_t, [_vp]),
    "pc_puncturer_positions": (_i, [_vp, _vp]),
    "pc_puncture_batch": (_i, [_vp, _vp, _vp, _i, ctypes.c_size_t, _vp]),
    "pc_depuncture_batch": (_i, [_vp, _vp, _vp, _i, ctypes.c_size_t, _vp]),
    "pc_puncture_packed_batch": (_i, [_vp, _vp, _vp, ctypes.c_size_t, _vp]),
    "pc_last_error": (ctypes.c_char_p, []),
    "pc_version": (ctypes.c_char_p, []),
    "pc_frozen_bits": (_i, [_i, _i, ctypes.c_float, ctypes.c_char_p, _vp]),
    "pc_crc_kind": (_i, [ctypes.c_uint, ctypes.c_char_p]),
    "pc_crc_bytes": (_i, [_i]),
    "pc_code_create": (_vp, [_i, _vp, _i, _i, _i]),
    "pc_code_destroy": (None, [_vp]),
    "pc_code_set_crc": (_i, [_vp, _i]),
    "pc_code_set_systematic": (_i, [_vp, _i]),
    "pc_code_N": (_i, [_vp]),
    "pc_code_K": (_i, [_vp]),
    "pc_code_info_bytes": (_i, [_vp]),
    "pc_code_program_length": (_i, [_vp]),
    "pc_code_program": (_i, [_vp, _vp, _i]),
    "pc_code_launch_info": (_i, [_vp, _vp, _vp, _vp]),
    "pc_decode_batch": (_i, [_vp, _vp, _i, _vp, _vp, _sz, _vp]),
    "pc_decode_batch_host": (_i, [_vp, _vp, _i, _vp, _vp, _sz]),
    "pc_decode_soft_batch": (_i, [_vp, _vp, _i, _vp, _vp, _sz, _vp]),
    "pc_decode_soft_batch_host": (_i, [_vp, _vp, _i, _vp, _vp, _sz]),
    "pc_decode_soft_batch_f32": (_i, [_vp, _vp, _i, _vp, _vp, _sz, _vp]),
    "pc_decode_soft_batch_f32_host": (_i, [_vp, _vp, _i, _vp, _vp, _sz]),
    "pc_decode_batch_f32_static": (_i, [_vp, _vp, _vp, _vp, _sz, _vp]),
    "pc_encode_batch": (_i, [_vp, _vp, _vp, _vp, _sz, _vp]),
    "pc_encode_batch_host": (_i, [_vp, _vp, _vp, _vp, _sz]),
    "pc_crc_generate_batch": (_i, [_i, _vp, _i, _sz, _vp]),
    "pc_crc_check_batch": (_i, [_i, _vp, _i, _vp, _sz, _vp]),
    "pc_crc_calculate": (_u64, [_i, _vp, _sz]),
    "pc_quantize_batch": (_i, [_i, _vp, _vp, _sz, _vp]),
    "pc_awgn_generate": (_i, [_vp, _u64, _u64, ctypes.c_float, ctypes.c_float, _i, _vp, _vp, _sz, _vp]),
    "pc_count_errors": (_i, [_vp, _vp, _vp, _i, _sz, _vp, _vp]),
    "pc_device_alloc": (_vp, [_sz]),
    "pc_device_free": (None, [_vp]),
    "pc_host_alloc": (_vp, [_sz]),
    "pc_host_free": (None, [_vp]),
    "pc_memcpy_h2d": (_i, [_vp, _vp, _sz]),
    "pc_memcpy_d2h": (_i, [_vp, _vp, _sz]),
    "pc_device_synchronize": (_i, []),
    "pc_launch_count": (ctypes.c_ulonglong, []),
}

_LIB = None


def lib():
    global _LIB
    if _LIB is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                f"{LIB_PATH} is missing: build it with `make -C polar-codes_b200/csrc` "
                "(or __graft_entry__.build()); this package has no CPU fallback")
        L = ctypes.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(L, name)
            fn.restype = res
            fn.argtypes = args
        _LIB = L
    return _LIB


class PolarError(RuntimeError):
    pass


def last_error():
    return (lib().pc_last_error() or b"").decode()


def check(status):
    if status != PC_OK:
        msg = last_error()
        if status == PC_ERR_ARG:
            raise ValueError(msg)
        raise PolarError(f"polar_b200 status {status}: {msg}")


def ptr(a):
    """void* of a numpy array, a torch tensor (host or device) or a raw integer address"""
    if a is None:
        return None
    if isinstance(a, int):
        return ctypes.c_void_p(a)
    if isinstance(a, np.ndarray):
        return a.ctypes.data_as(ctypes.c_void_p)
    if hasattr(a, "data_ptr"):
        return ctypes.c_void_p(a.data_ptr())
    raise TypeError(f"cannot take the address of {type(a)}")
