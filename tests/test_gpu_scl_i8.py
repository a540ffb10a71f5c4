"""GPU parity: CRC-aided SCL int8 kernel against the unmodified reference's SclFipChar (oracle/_ref) and our C
restatement (oracle/pc_oracle_scl.c) on identical LLRs. Bar: bit-exact info bytes and verdict -- this includes the
survivor order and every tie of the reference's swap-based selection sorts."""
import numpy as np
import pytest

from conftest import awgn_llr_i8

pytestmark = pytest.mark.gpu


def _tx_llrs(ref, rng, N, K, fz, crc_bits, crc_type, B):
    info = rng.integers(0, 256, (B, (K + 7) // 8), dtype=np.uint8)
    if K % 8:
        info[:, -1] &= (0xFF00 >> (K % 8)) & 0xFF
    code, _ = ref.encode_batch(N, fz, info, crc_bits=crc_bits, crc_type=crc_type)
    cw = np.unpackbits(code, axis=1, bitorder="big")[:, :N]
    return np.concatenate([awgn_llr_i8(rng, B, N, 1.0, K / N, 10.0, codewords=cw),
                           awgn_llr_i8(rng, B, N, 2.5, K / N, 10.0, codewords=cw),
                           awgn_llr_i8(rng, B, N, 0.0, K / N, 40.0, codewords=cw),
                           rng.integers(-128, 128, (max(B // 4, 8), N), dtype=np.int8),
                           rng.choice(np.array([-128, -127, 127, 0, 1, -1], np.int8), (max(B // 4, 8), N))])


CASES = [(1024, 512, 8, 16, "crcnr"), (8, 4, 4, 8, "crc"), (64, 40, 4, 8, "crc"), (512, 128, 8, 24, "crc"),
         (256, 128, 2, 0, "crc"), (128, 64, 16, 8, "crc"), (1024, 768, 8, 32, "crc"), (32, 16, 8, 6, "crc"),
         (1024, 512, 32, 16, "crc"), (1024, 512, 4, 11, "crc"), (2048, 1024, 8, 16, "crc"), (256, 200, 3, 8, "crc"),
         (64, 32, 6, 8, "crc"), (16, 8, 32, 8, "crc")]


@pytest.mark.parametrize("N,K,L,crc_bits,crc_type", CASES)
def test_scl_parity(pcb, ref, N, K, L, crc_bits, crc_type):
    rng = np.random.default_rng(N + 3 * K + L)
    fz = ref.frozen_bits(N, K, 0.0)
    B = 256 if L <= 8 else 64
    llr = _tx_llrs(ref, rng, N, K, fz, crc_bits, crc_type, B)
    dec = pcb.PolarDecoder(N, L, list(fz), "char")
    dec.setErrorDetection(crc_bits, crc_type)
    got, ok = dec.decode_batch(llr, return_ok=True)
    exp, eok = ref.decode_batch(N, L, fz, "char", llr, crc_bits=crc_bits, crc_type=crc_type)
    bad = (got != exp).any(axis=1)
    assert not bad.any(), f"{bad.sum()} of {len(bad)} frames differ (first {np.flatnonzero(bad)[:5]})"
    assert np.array_equal(ok, eok)


def test_scl_kat_n8(pcb):
    """reference test/polarcode/decodingtest.cpp:1137-1181 (testListDecoder)"""
    dec = pcb.PolarDecoder(8, 4, [0, 1, 2, 4], "char")
    out = dec.decode_vector(np.array([-5, -6, -4, 1, -4, -5, -7, 2], dtype=np.int8))
    assert (out[0] & 0xF0) == 0xF0


def test_scl_non_systematic(pcb, ref):
    N, K, L = 1024, 512, 8
    rng = np.random.default_rng(11)
    fz = ref.frozen_bits(N, K, 0.0)
    llr = awgn_llr_i8(rng, 300, N, 2.0, K / N)
    dec = pcb.PolarDecoder(N, L, list(fz), "char")
    dec.setSystematic(False)
    got, ok = dec.decode_batch(llr, return_ok=True)
    exp, eok = ref.decode_batch(N, L, fz, "char", llr, systematic=False)
    assert np.array_equal(got, exp) and np.array_equal(ok, eok)


def test_scl_vs_restatement_and_ragged(pcb, orc):
    N, K, L = 512, 256, 8
    rng = np.random.default_rng(12)
    fz = orc.frozen_bits(N, K, 0.0)
    dec = pcb.PolarDecoder(N, L, list(fz), "char")
    dec.setErrorDetection(16, "crc")
    for B in (1, 15, 17, 333):
        llr = awgn_llr_i8(rng, B, N, 1.5, K / N)
        got, ok = dec.decode_batch(llr, return_ok=True)
        exp, eok = orc.scl_i8_decode(N, fz, L, llr, crc=16)
        assert np.array_equal(got, exp) and np.array_equal(ok, eok)
