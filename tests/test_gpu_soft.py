"""Soft outputs of the int8 and fp32 Fast-SSC decoders: Decoder::getSoftCodeword / getSoftInformation
(reference src/polarcode/decoding/decoder.cpp:147-152) byte for byte against the unmodified reference, plus the
reference's own soft-output KATs (test/polarcode/decodingtest.cpp:83-156: REP / DoubleREP saturation)."""
import numpy as np
import pytest

from conftest import awgn_llr_i8

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("N,K", [(1024, 512), (64, 40), (8, 4), (16, 8), (32, 16), (128, 64), (256, 64), (512, 128),
                                 (2048, 1800), (4096, 3072), (1024, 1023), (1024, 1), (16384, 8192)])
def test_soft_outputs_match_reference(pcb, ref, N, K):
    rng = np.random.default_rng(N + K)
    fz = ref.frozen_bits(N, K, 0.0)
    dec = pcb.PolarDecoder(N, 1, list(fz), "char")
    B = 4096 if N <= 1024 else (512 if N <= 4096 else 64)
    batches = [awgn_llr_i8(rng, B, N, 2.0, K / N, 10.0), awgn_llr_i8(rng, B // 4, N, 0.0, K / N, 45.0),
               rng.integers(-128, 128, (B // 4, N), dtype=np.int8),
               rng.choice(np.array([-128, -127, 127, 0, 1, -1], np.int8), (B // 4, N)),
               np.stack([np.full(N, v, np.int8) for v in (127, -127, -128, 0, 1, -1)])]
    for llr in batches:
        sc, si = dec.decode_soft_batch(llr)
        esc, esi = ref.decode_soft(N, fz, llr)
        assert np.array_equal(sc, esc), f"soft codeword differs in {(sc != esc).any(axis=1).sum()} frames"
        assert np.array_equal(si, esi)
        # the decisions of the bit-packed kernel are the signs of these bytes (systematic: info = codeword positions)
        hard = dec.decode_batch(llr)
        assert np.array_equal(np.packbits(si.view(np.uint8) >> 7, axis=1), hard)


def test_rep_and_double_rep_soft_kats(pcb):
    """decodingtest.cpp:83-111 (REP, N = 32 / 64) and :118-156 (DoubleREP): the soft output is the saturated sum"""
    for N in (32, 64):
        fz = list(range(N - 1))  # one information bit: a REP node
        dec = pcb.PolarDecoder(N, 1, fz, "char")
        for fill, exp in [(1, N), (3, min(3 * N, 127)), (-3, max(-3 * N, -128)), (-1, -N)]:
            sc, _ = dec.decode_soft_batch(np.full((1, N), fill, np.int8))
            assert (sc[0] == exp).all()
    N = 64
    dec = pcb.PolarDecoder(N, 1, list(range(N - 2)), "char")  # two information bits at the end: DoubleREP
    llr = np.zeros((1, N), np.int8)
    llr[0, 0::2], llr[0, 1::2] = 3, -2
    sc, _ = dec.decode_soft_batch(llr)
    assert (sc[0, 0::2] == min(3 * 32, 127)).all() and (sc[0, 1::2] == max(-2 * 32, -128)).all()


def test_soft_device_tensors_and_float_input(pcb, ref):
    import torch
    N, K = 1024, 512
    rng = np.random.default_rng(3)
    fz = ref.frozen_bits(N, K, 0.0)
    dec = pcb.PolarDecoder(N, 1, list(fz), "char")
    y = (10.0 * (1.0 + 0.7 * rng.standard_normal((2000, N)))).astype(np.float32)
    sc, si = dec.decode_soft_batch(torch.from_numpy(y).cuda())
    esc, esi = ref.decode_soft(N, fz, ref.quantize(N, y))
    assert np.array_equal(sc.cpu().numpy(), esc) and np.array_equal(si.cpu().numpy(), esi)


def test_soft_refused_for_other_kinds(pcb, ref):
    fz = ref.frozen_bits(256, 128, 0.0)
    for kind, L in (("float", 4), ("char", 4), ("mixed", 4)):
        dec = pcb.PolarDecoder(256, L, list(fz), kind)
        with pytest.raises(Exception, match="soft outputs"):
            dec.decode_soft_batch(np.zeros((2, 256), np.int8))


@pytest.mark.parametrize("N,K", [(8, 4), (32, 16), (64, 40), (128, 100), (256, 128), (512, 256), (1024, 512), (1024, 900), (1024, 100),
                                 (2048, 1024), (4096, 2048)])
def test_float_soft_outputs_match_reference(pcb, ref, N, K):
    """getSoftCodeword / getSoftInformation of the 'float' Fast-SSC decoder: the float bit container as 32-bit patterns"""
    rng = np.random.default_rng(11 * N + K)
    fz = ref.frozen_bits(N, K, 0.0)
    dec = pcb.PolarDecoder(N, 1, list(fz), "float")
    llr = np.concatenate([(a * (1.0 + s * rng.standard_normal((300, N)))).astype(np.float32) for a, s in ((2.0, 1.0), (8.0, 0.7), (1.0, 3.0))] +
                         [rng.choice(np.array([0.0, -0.0, 1.0, -1.0, 2.5], np.float32), (100, N))])
    sc, si = dec.decode_soft_batch(llr)
    esc, esi = ref.decode_soft_f32(N, fz, llr)
    bad = (sc.view(np.uint32) != esc.view(np.uint32)).any(axis=1)
    assert not bad.any(), f"soft codeword differs in {bad.sum()} frames"
    assert np.array_equal(si.view(np.uint32), esi.view(np.uint32))
    # the hard decisions are the signs of that container
    out = dec.decode_batch(llr)
    exp, _ = ref.decode_batch(N, 1, fz, "float", llr, crc_bits=8)
    assert np.array_equal(out, exp)


def test_float_soft_device_tensors_int8_input_and_c2_batch(pcb, ref):
    import torch
    N, K = 1024, 512
    fz = ref.frozen_bits(N, K, 0.0)
    dec = pcb.PolarDecoder(N, 1, list(fz), "float")
    rng = np.random.default_rng(5)
    y = (4.0 * 10 ** 0.2 * (1.0 + 0.79 * rng.standard_normal((4096, N)))).astype(np.float32)
    sc, si = dec.decode_soft_batch(torch.from_numpy(y).cuda())
    esc, esi = ref.decode_soft_f32(N, fz, y)
    assert np.array_equal(sc.cpu().numpy().view(np.uint32), esc.view(np.uint32))
    assert np.array_equal(si.cpu().numpy().view(np.uint32), esi.view(np.uint32))
    q = np.clip(np.rint(y[:512] * 4), -128, 127).astype(np.int8)  # FloatContainer::insertLlr(const char*): int8 -> float
    sc8, _ = dec.decode_soft_batch(q)
    esc8, _ = ref.decode_soft_f32(N, fz, q.astype(np.float32))
    assert np.array_equal(sc8.view(np.uint32), esc8.view(np.uint32))
