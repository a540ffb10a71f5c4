"""GPU parity of the fp32 decoders against the unmodified reference (FastSscAvxFloat, SclAvxFloat, AdaptiveFloat) on
identical float LLRs. north_star bar: identical hard decisions on >= 99.99 % of frames; the kernels restate the
reference's lane structure (order of float additions, tie rules), so the tests assert 100 % on these seeds and the
tolerance written here is the north_star one."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

MIN_FRAME_AGREEMENT = 0.9999


def _llrs(rng, B, N, sigma=0.8):
    y = (1.0 + sigma * rng.standard_normal((B, N))).astype(np.float32)
    return [(4.0 * y).astype(np.float32),                                  # pcsim float scaling
            np.clip(np.rint(10 * y), -128, 127).astype(np.float32),        # integer valued: many exact ties
            np.rint(2 * y).astype(np.float32)]                             # coarse: zeros and ties everywhere


def _agree(got, exp):
    return 1.0 - (got != exp).any(axis=1).mean()


FSSC = [(1024, 512), (64, 40), (8, 4), (16, 8), (32, 16), (128, 64), (256, 64), (512, 128), (2048, 1800), (1024, 256),
        (1024, 768), (4096, 2048), (64, 8), (64, 56), (16, 12), (16, 4), (8, 6), (8, 2), (32, 5), (32, 27)]


@pytest.mark.parametrize("N,K", FSSC)
def test_fastssc_float_parity(pcb, ref, N, K):
    rng = np.random.default_rng(N * 3 + K)
    for dsnr in (0.0, 1.0, 3.0):
        fz = ref.frozen_bits(N, K, dsnr)
        dec = pcb.PolarDecoder(N, 1, list(fz), "float")
        for llr in _llrs(rng, 400 if N <= 1024 else 100, N):
            got, ok = dec.decode_batch(llr, return_ok=True)
            exp, eok = ref.decode_batch(N, 1, fz, "float", llr)
            assert _agree(got, exp) >= MIN_FRAME_AGREEMENT
            assert np.array_equal(got, exp) and np.array_equal(ok, eok)


def test_fastssc_float_int8_input_and_nonsystematic(pcb, ref):
    N, K = 1024, 512
    rng = np.random.default_rng(1)
    fz = ref.frozen_bits(N, K, 0.0)
    dec = pcb.PolarDecoder(N, 1, list(fz), "float")
    llr8 = np.clip(np.rint(10 * (1 + 0.7 * rng.standard_normal((500, N)))), -128, 127).astype(np.int8)
    got = dec.decode_batch(llr8)
    exp, _ = ref.decode_batch(N, 1, fz, "float", llr8)  # FloatContainer::insertLlr(const char*)
    assert np.array_equal(got, exp)
    dec.setSystematic(False)
    llr = _llrs(rng, 300, N)[0]
    got, ok = dec.decode_batch(llr, return_ok=True)
    exp, eok = ref.decode_batch(N, 1, fz, "float", llr, systematic=False)
    assert np.array_equal(got, exp) and np.array_equal(ok, eok)


def test_float_kat_n8(pcb):
    """reference test/polarcode/decodingtest.cpp:1137-1181 pattern on the float Fast-SSC decoder"""
    dec = pcb.PolarDecoder(8, 1, [0, 1, 2, 4], "float")
    out = dec.decode_vector(np.array([-5, -6, -4, 1, -4, -5, -7, 2], dtype=np.float32))
    assert (out[0] & 0xF0) == 0xF0


# ---- SCL fp32 (SclAvxFloat) and the adaptive 'mixed' decoder (AdaptiveFloat) -------------------------------------------
def _tx_float_llrs(ref, rng, N, K, fz, crc_bits, crc_type, B):
    """transmitted codewords (so that the detector passes on some list path) at three noise levels, as float LLRs in
    pcsim's scaling, integer-valued (exact metric ties) and very coarse (zeros and ties everywhere)"""
    info = rng.integers(0, 256, (B, (K + 7) // 8), dtype=np.uint8)
    if K % 8:
        info[:, -1] &= (0xFF00 >> (K % 8)) & 0xFF
    code, _ = ref.encode_batch(N, fz, info, crc_bits=crc_bits, crc_type=crc_type)
    x = 1.0 - 2.0 * np.unpackbits(code, axis=1, bitorder="big")[:, :N].astype(np.float32)
    out = []
    for sigma, scale, rint in ((0.7, 4.0, False), (0.9, 10.0, True), (1.0, 2.0, True), (0.5, 3.0, False), (1.0, 4.0, False)):
        y = (x + sigma * rng.standard_normal((B, N))).astype(np.float32) * np.float32(scale)
        out.append(np.rint(y).astype(np.float32) if rint else y.astype(np.float32))
    return np.concatenate(out)


SCLF = [(64, 40, 4, 8, "crc"), (512, 128, 8, 24, "crc"), (1024, 512, 8, 16, "crc"), (8, 4, 4, 8, "crc"),
        (1024, 512, 32, 16, "crc"), (256, 128, 2, 0, "crc"), (128, 64, 16, 8, "crc"), (32, 16, 8, 6, "crc"),
        (2048, 1024, 4, 16, "crcnr"), (256, 200, 3, 8, "crc"), (16, 8, 32, 8, "crc"), (1024, 768, 8, 32, "crc"),
        (4096, 2048, 2, 11, "crc")]


@pytest.mark.parametrize("N,K,L,crc_bits,crc_type", SCLF)
def test_scl_float_parity(pcb, ref, N, K, L, crc_bits, crc_type):
    rng = np.random.default_rng(7 * N + K + L)
    fz = ref.frozen_bits(N, K, 0.0)
    B = 128 if L <= 8 and N <= 1024 else 32
    llr = _tx_float_llrs(ref, rng, N, K, fz, crc_bits, crc_type, B)
    dec = pcb.PolarDecoder(N, L, list(fz), "float")
    dec.setErrorDetection(crc_bits, crc_type)
    got, ok = dec.decode_batch(llr, return_ok=True)
    # fresh=True: a new reference object per frame; a long-lived SclAvxFloat carries its best path metric from frame to
    # frame (tests/test_oracle.py::test_reference_float_list_decoder_keeps_metric_across_frames)
    exp, eok = ref.decode_batch(N, L, fz, "float", llr, crc_bits=crc_bits, crc_type=crc_type, fresh=True, threads=8)
    assert _agree(got, exp) >= MIN_FRAME_AGREEMENT
    bad = (got != exp).any(axis=1)
    assert not bad.any(), f"{bad.sum()} of {len(bad)} frames differ (first {np.flatnonzero(bad)[:5]})"
    assert np.array_equal(ok, eok)


@pytest.mark.parametrize("N,K,L,dsnr", [(64, 40, 4, 1.0), (1024, 512, 8, 0.0), (512, 128, 8, 0.0), (256, 128, 1, 0.0)])
def test_adaptive_mixed_parity(pcb, ref, N, K, L, dsnr):
    """configs[0]: README pypolar case N=64 K=40 (BB, 1.0 dB), L=4 'mixed'; AdaptiveFloat adaptive_float.cpp:33-45"""
    rng = np.random.default_rng(N + K + L)
    fz = ref.frozen_bits(N, K, dsnr)
    llr = _tx_float_llrs(ref, rng, N, K, fz, 8, "crc", 256)
    dec = pcb.PolarDecoder(N, L, list(fz), "mixed")
    got, ok = dec.decode_batch(llr, return_ok=True)
    exp, eok = ref.decode_batch(N, L, fz, "mixed", llr, fresh=True, threads=8)
    assert N not in (64, 1024) or 0 < ok.mean() < 1  # both passes are exercised
    assert np.array_equal(got, exp) and np.array_equal(ok, eok)
    # ragged batch sizes and the non-systematic flavour
    for B in (1, 33):
        got, ok = dec.decode_batch(llr[:B], return_ok=True)
        assert np.array_equal(got, exp[:B]) and np.array_equal(ok, eok[:B])
    if N >= 256:
        dec.setSystematic(False)
        got, ok = dec.decode_batch(llr, return_ok=True)
        exp, eok = ref.decode_batch(N, L, fz, "mixed", llr, systematic=False, fresh=True, threads=8)
        assert np.array_equal(got, exp) and np.array_equal(ok, eok)


def test_scl_float_kat_n8_and_int8_input(pcb, ref):
    """reference test/polarcode/decodingtest.cpp:1137-1181 (testListDecoder, float flavour)"""
    dec = pcb.PolarDecoder(8, 4, [0, 1, 2, 4], "float")
    out = dec.decode_vector(np.array([-5, -6, -4, 1, -4, -5, -7, 2], dtype=np.float32))
    assert (out[0] & 0xF0) == 0xF0
    N, K, L = 512, 256, 4
    rng = np.random.default_rng(5)
    fz = ref.frozen_bits(N, K, 0.0)
    llr8 = np.clip(np.rint(10 * (1 + 0.8 * rng.standard_normal((200, N)))), -128, 127).astype(np.int8)
    got, ok = pcb.PolarDecoder(N, L, list(fz), "float").decode_batch(llr8, return_ok=True)
    exp, eok = ref.decode_batch(N, L, fz, "float", llr8, fresh=True)
    assert np.array_equal(got, exp) and np.array_equal(ok, eok)
