"""GPU parity: Fast-SSC int8 batch kernel (through the C-ABI) against the oracle on identical LLRs.

Bar: bit-exact info bytes AND detector verdict (reference src/polarcode/decoding/fastssc_fip_char.cpp:616-630).
Oracle = the unmodified reference (oracle/_ref) where built, else our C restatement (oracle/pc_oracle.c).
"""
import numpy as np
import pytest

from conftest import awgn_llr_i8

pytestmark = pytest.mark.gpu


def _oracle_decode(ref, orc, N, fz, llr, crc_bits=8, crc_type="crc", systematic=True):
    import refshim
    if refshim.available():
        return ref.decode_batch(N, 1, fz, "char", llr, crc_bits=crc_bits, crc_type=crc_type, systematic=systematic)
    out = orc.fastssc_i8_decode(N, fz, llr, systematic=systematic)
    ok = orc.crc_check(orc.crc_kind(crc_bits, crc_type), out) if crc_bits else np.ones(len(out), np.uint8)
    return out, ok


def _dec(pcb, N, fz, crc_bits=8, crc_type="crc", systematic=True):
    d = pcb.PolarDecoder(N, 1, list(fz), "char")
    d.setErrorDetection(crc_bits, crc_type)
    if not systematic:
        d.setSystematic(False)
    return d


CODES = [(1024, 512), (64, 40), (8, 4), (8, 7), (16, 8), (32, 16), (128, 64), (256, 64), (512, 128), (2048, 1800),
         (4096, 3072), (1024, 1023), (1024, 1), (16384, 8192)]


@pytest.mark.parametrize("N,K", CODES)
def test_awgn_parity(pcb, ref, orc, N, K):
    rng = np.random.default_rng(N * 7 + K)
    fz = ref.frozen_bits(N, K, 0.0)
    dec = _dec(pcb, N, fz)
    B = 4096 if N <= 4096 else 600
    for ebn0, amp in [(0.0, 10.0), (2.0, 10.0), (4.0, 10.0), (1.0, 45.0), (3.0, 2.0)]:
        llr = awgn_llr_i8(rng, B, N, ebn0, K / N, amp)
        got, ok = dec.decode_batch(llr, return_ok=True)
        exp, eok = _oracle_decode(ref, orc, N, fz, llr)
        assert np.array_equal(got, exp), f"info mismatch in {(got != exp).any(axis=1).sum()} frames"
        assert np.array_equal(ok, eok)


@pytest.mark.parametrize("N,K", [(1024, 512), (64, 40), (256, 200), (16384, 8192)])
def test_adversarial_parity(pcb, ref, orc, N, K):
    """saturation floods: +-127, -128, zeros, uniform random bytes"""
    rng = np.random.default_rng(5)
    fz = ref.frozen_bits(N, K, 0.0)
    dec = _dec(pcb, N, fz)
    B = 512 if N <= 4096 else 64
    cases = [np.full((4, N), v, np.int8) for v in (127, -127, -128, 0, 1, -1)]
    cases.append(rng.integers(-128, 128, (B, N), dtype=np.int8))
    cases.append(rng.choice(np.array([-128, -127, 127, 0], np.int8), (B, N)))
    cases.append(rng.choice(np.array([-128, 127], np.int8), (B, N)))
    cases.append(rng.choice(np.array([-1, 0, 1], np.int8), (B, N)))
    llr = np.concatenate(cases)
    got, ok = dec.decode_batch(llr, return_ok=True)
    exp, eok = _oracle_decode(ref, orc, N, fz, llr)
    assert np.array_equal(got, exp)
    assert np.array_equal(ok, eok)


@pytest.mark.parametrize("crc_bits,crc_type", [(0, "crc"), (6, "crc"), (8, "crc"), (11, "crc"), (16, "crc"),
                                               (16, "crcnr"), (24, "crc"), (32, "crc")])
def test_detector_verdict(pcb, ref, orc, crc_bits, crc_type):
    N, K = 1024, 512
    rng = np.random.default_rng(crc_bits)
    fz = ref.frozen_bits(N, K, 0.0)
    dec = _dec(pcb, N, fz, crc_bits, crc_type)
    # transmit valid CRC-carrying codewords so that both verdicts occur
    info = rng.integers(0, 256, (2048, K // 8), dtype=np.uint8)
    code, _ = ref.encode_batch(N, fz, info, crc_bits=crc_bits, crc_type=crc_type)
    cw = np.unpackbits(code, axis=1, bitorder="big")
    llr = awgn_llr_i8(rng, 2048, N, 2.5, K / N, 10.0, codewords=cw)
    got, ok = dec.decode_batch(llr, return_ok=True)
    exp, eok = _oracle_decode(ref, orc, N, fz, llr, crc_bits, crc_type)
    assert np.array_equal(got, exp)
    assert np.array_equal(ok, eok)
    if crc_bits:
        assert 0 < ok.sum() < len(ok) or ok.all()


# the reference's non-systematic decode does not round-trip for N < 256 (PackedContainer mFakeSize offset,
# bitcontainer.cpp:631-713; verified with oracle/_ref) -- parity is asserted for N >= 256 only
@pytest.mark.parametrize("N,K", [(1024, 512), (256, 128), (4096, 3072)])
def test_non_systematic(pcb, ref, orc, N, K):
    rng = np.random.default_rng(3)
    fz = ref.frozen_bits(N, K, 0.0)
    dec = _dec(pcb, N, fz, systematic=False)
    llr = awgn_llr_i8(rng, 1024, N, 2.0, K / N)
    got, ok = dec.decode_batch(llr, return_ok=True)
    exp, eok = _oracle_decode(ref, orc, N, fz, llr, systematic=False)
    assert np.array_equal(got, exp)
    assert np.array_equal(ok, eok)


def test_kat_n8(pcb):
    """reference test/polarcode/decodingtest.cpp:1031-1060 (testFipShort)"""
    dec = pcb.PolarDecoder(8, 1, [0, 1, 2, 4], "char")
    out = dec.decode_vector(np.array([-5, -6, -4, 1, -4, -5, -7, 2], dtype=np.int8))
    assert (out[0] & 0xF0) == 0xF0


def test_ragged_and_empty(pcb, ref, orc):
    N, K = 1024, 512
    fz = ref.frozen_bits(N, K, 0.0)
    dec = _dec(pcb, N, fz)
    assert dec.decode_batch(np.zeros((0, N), np.int8)).shape == (0, K // 8)
    rng = np.random.default_rng(9)
    for B in (1, 31, 33, 191, 193, 148 * 192 + 7):
        llr = awgn_llr_i8(rng, B, N, 1.0, 0.5)
        got = dec.decode_batch(llr)
        exp, _ = _oracle_decode(ref, orc, N, fz, llr)
        assert np.array_equal(got, exp)
    with pytest.raises(RuntimeError):
        dec.decode_vector(np.zeros(N - 1, np.int8))


def test_device_tensor_path_full_size(pcb):
    """BASELINE size (2^20 frames): size-independent property -- the all-zero codeword at high SNR decodes to zeros,
    a CRC-8-valid word, and results are identical between two launches (determinism)"""
    import torch
    N, K = 1024, 512
    fz = pcb.frozen_bits(N, K, 0.0)
    dec = pcb.PolarDecoder(N, 1, fz, "char")
    g = torch.Generator(device="cuda").manual_seed(1)
    B = 1 << 18
    y = 1.0 + 0.45 * torch.randn((B, N), device="cuda", generator=g)
    llr = torch.clamp(torch.round(10.0 * y), -128, 127).to(torch.int8)
    out, ok = dec.decode_batch(llr, return_ok=True)
    out2 = dec.decode_batch(llr)
    torch.cuda.synchronize()
    assert torch.equal(out, out2)
    assert (out == 0).all(dim=1).float().mean().item() > 0.999
    assert ok.float().mean().item() > 0.999
