"""pc_decode_batch_f32_static (include/polar_b200.h): the allocation-free fp32 Fast-SSC decode whose scratch is a file-scope
device array. Called through the C-ABI twice in a row and compared bit for bit with pc_decode_batch on the same LLRs (the
tolerance of every fp32 Fast-SSC test here: identical info bytes and verdicts). Collected last on purpose (file name)."""
import ctypes

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _frames(rng, B, N, K, ebn0_db, scale):
    sigma = 1.0 / np.sqrt(2.0 * (K / N) * 10.0 ** (ebn0_db / 10.0))
    y = 1.0 + sigma * rng.standard_normal((B, N))  # all-zero codeword, BPSK 0 -> +1
    return np.ascontiguousarray((scale * y).astype(np.float32))


@pytest.mark.parametrize("N,K,B", [(1024, 512, 20000), (256, 128, 3000), (4096, 2048, 1500), (64, 40, 777)])
def test_static_scratch_decode_matches_decode_batch(pcb, N, K, B):
    capi = pcb.capi
    L = capi.lib()
    vp = ctypes.c_void_p
    fz = pcb.frozen_bits(N, K, 0.0)
    dec = pcb.PolarDecoder(N, 1, fz, "float")
    dec.setErrorDetection(8, "crc")
    kb = (K + 7) // 8
    d_llr = vp(L.pc_device_alloc(4 * B * N))
    d_info, d_ok = vp(L.pc_device_alloc(B * kb)), vp(L.pc_device_alloc(B))
    d_info2, d_ok2 = vp(L.pc_device_alloc(B * kb)), vp(L.pc_device_alloc(B))
    assert d_llr.value and d_info.value and d_ok.value and d_info2.value and d_ok2.value
    try:
        rng = np.random.default_rng(N + B)
        for call, (ebn0, scale) in enumerate(((2.0, 4.0 * 10 ** 0.2), (1.0, 1.0))):  # two calls in a row, different data
            llr = _frames(rng, B, N, K, ebn0, scale)
            capi.check(L.pc_memcpy_h2d(d_llr, capi.ptr(llr), llr.nbytes))
            capi.check(L.pc_decode_batch(dec._h, d_llr, 1, d_info, d_ok, B, None))
            capi.check(L.pc_decode_batch_f32_static(dec._h, d_llr, d_info2, d_ok2, B, None))
            capi.check(L.pc_device_synchronize())
            a, ok_a = np.empty((B, kb), np.uint8), np.empty(B, np.uint8)
            b, ok_b = np.empty((B, kb), np.uint8), np.empty(B, np.uint8)
            capi.check(L.pc_memcpy_d2h(capi.ptr(a), d_info, a.nbytes))
            capi.check(L.pc_memcpy_d2h(capi.ptr(ok_a), d_ok, B))
            capi.check(L.pc_memcpy_d2h(capi.ptr(b), d_info2, b.nbytes))
            capi.check(L.pc_memcpy_d2h(capi.ptr(ok_b), d_ok2, B))
            assert np.array_equal(a, b), f"call {call}: {(a != b).any(axis=1).sum()} of {B} frames differ"
            assert np.array_equal(ok_a, ok_b)
            assert 0 < ok_a.sum() < B or N == 64  # both verdicts occur: the comparison is not vacuous
    finally:
        for d in (d_llr, d_info, d_ok, d_info2, d_ok2):
            L.pc_device_free(d)


def test_static_scratch_limits(pcb):
    """beyond the stated maximum (N <= 4096) and for other decoder kinds the call is refused with a clear error"""
    capi = pcb.capi
    L = capi.lib()
    vp = ctypes.c_void_p
    d = vp(L.pc_device_alloc(1 << 20))
    try:
        big = pcb.PolarDecoder(8192, 1, pcb.frozen_bits(8192, 4096, 0.0), "float")
        assert L.pc_decode_batch_f32_static(big._h, d, d, None, 1, None) == capi.PC_ERR_UNSUPPORTED
        assert "4096" in capi.last_error()
        other = pcb.PolarDecoder(1024, 1, pcb.frozen_bits(1024, 512, 0.0), "char")
        assert L.pc_decode_batch_f32_static(other._h, d, d, None, 1, None) == capi.PC_ERR_UNSUPPORTED
        assert "float" in capi.last_error()
    finally:
        L.pc_device_free(d)
