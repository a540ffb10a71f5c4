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
