"""CPU tests: pin the oracle (oracle/pc_oracle.c, our scalar C restatement) against
 (1) the reference's own known-answer tests (file:line cited per test),
 (2) golden vectors generated from the unmodified reference (tests/golden/make_golden.py),
 (3) the unmodified reference itself where oracle/_ref is built (differential, random + adversarial inputs)."""
import os

import numpy as np
import pytest

GOLD = np.load(os.path.join(os.path.dirname(__file__), "golden", "reference_vectors.npz"))


def test_construction_kats(orc):
    """reference test/polarcode/constructiontest.cpp:22-78"""
    assert orc.frozen_bits(8, 4, 0.0).tolist() == [0, 1, 2, 4]
    for key in GOLD.files:
        if key.startswith("frozen_"):
            _, t, N, K, d = key.split("_")
            assert np.array_equal(orc.frozen_bits(int(N), int(K), float(d), t), GOLD[key]), key


def test_f_g_primitives_exhaustive(orc):
    """include/polarcode/decoding/fip_char.h:44-73; KATs test/polarcode/decodingtest.cpp:921-933 (SSE twin literals)"""
    a = np.repeat(np.arange(-128, 128, dtype=np.int16), 256).astype(np.int8)
    b = np.tile(np.arange(-128, 128, dtype=np.int16), 256).astype(np.int8)
    f = orc.f_i8(a, b).astype(np.int32)
    ai, bi = a.astype(np.int32), b.astype(np.int32)
    mag = np.maximum(np.minimum(np.minimum(np.abs(ai), 127), np.minimum(np.abs(bi), 127)), 1)
    sign = np.where((ai < 0) ^ (bi < 0), -1, 1)
    assert np.array_equal(f, sign * mag)
    assert (f != 0).all() and (f != -128).all()
    for bit in (0, -128, 127, -1):
        g = orc.g_i8(a, b, np.full(a.shape, bit, np.int8)).astype(np.int32)
        exp = np.clip(bi - ai, -128, 127) if bit < 0 else np.clip(bi + ai, -128, 127)
        assert np.array_equal(g, exp)


def test_decoder_kat_n8(orc):
    """test/polarcode/decodingtest.cpp:1031-1060"""
    out = orc.fastssc_i8_decode(8, np.array([0, 1, 2, 4], np.uint32), np.array([[-5, -6, -4, 1, -4, -5, -7, 2]], np.int8))
    assert (out[0, 0] & 0xF0) == 0xF0


@pytest.mark.parametrize("N,K", [(1024, 512), (64, 40), (8, 4), (256, 64), (2048, 1800), (16384, 8192)])
def test_decoder_golden(orc, N, K):
    fz, llr = GOLD[f"dec_{N}_{K}_frozen"], GOLD[f"dec_{N}_{K}_llr"]
    out = orc.fastssc_i8_decode(N, fz, llr)
    assert np.array_equal(out, GOLD[f"dec_{N}_{K}_info"])
    assert np.array_equal(orc.crc_check(8, out), GOLD[f"dec_{N}_{K}_ok"])


def test_decoder_golden_nonsystematic(orc):
    fz = GOLD["dec_1024_512_frozen"]
    out = orc.fastssc_i8_decode(1024, fz, GOLD["decns_1024_512_llr"], systematic=False)
    assert np.array_equal(out, GOLD["decns_1024_512_info"])


def test_rep_soft_output_saturation(orc):
    """test/polarcode/decodingtest.cpp:83-111: REP soft output = clamp(sum), order-normative saturating tree"""
    N = 64
    fz = np.arange(N - 1, dtype=np.uint32)
    for fill, exp in [(1, 64), (3, 127), (-3, -128), (-1, -64)]:
        _, soft = orc.fastssc_i8_decode(N, fz, np.full((1, N), fill, np.int8), soft=True)
        assert (soft[0] == exp).all()


@pytest.mark.parametrize("N,K", [(1024, 512), (64, 40), (8, 4), (256, 64), (2048, 1800)])
def test_soft_codeword_matches_reference(orc, ref, N, K):
    """Decoder::getSoftCodeword (decoder.cpp:147-152): the restatement's root bit bytes against the reference's"""
    rng = np.random.default_rng(N)
    fz = ref.frozen_bits(N, K, 0.0)
    llr = np.concatenate([rng.integers(-128, 128, (64, N), dtype=np.int8),
                          np.clip(np.rint(10 * (1 + 0.8 * rng.standard_normal((192, N)))), -128, 127).astype(np.int8)])
    _, soft = orc.fastssc_i8_decode(N, fz, llr, soft=True)
    esc, esi = ref.decode_soft(N, fz, llr)
    assert np.array_equal(soft, esc)
    info_pos = np.setdiff1d(np.arange(N), fz)
    assert np.array_equal(esi, esc[:, info_pos])


@pytest.mark.parametrize("N,K", [(1024, 512), (256, 128), (4096, 3072), (16384, 8192)])
def test_encoder_golden(orc, N, K):
    fz = orc.frozen_bits(N, K, 0.0)
    for syst in (1, 0):
        code = orc.encode(N, fz, GOLD[f"enc_{N}_{K}_{syst}_info"], systematic=bool(syst))
        assert np.array_equal(code, GOLD[f"enc_{N}_{K}_{syst}_code"])


def test_encoder_kats(orc):
    """test/polarcode/encodingtest.cpp:26-90"""
    for n in range(3, 11):
        N = 1 << n
        u = np.zeros((1, N // 8), np.uint8)
        u[0, -1] = 1
        assert (orc.encode(N, np.zeros(0, np.uint32), u, systematic=False) == 0xFF).all()
    assert orc.encode(16, orc.frozen_bits(16, 8, 0.0), np.array([[0xFF]], np.uint8)).tolist() == [[0xFF, 0xFF]]


def test_encoder_systematic_property(orc):
    """test/polarcode/polarcodetest.cpp:45-100: the info bits reappear at the non-frozen positions"""
    N, K = 4096, 3072
    fz = orc.frozen_bits(N, K, 0.0)
    rng = np.random.default_rng(1001)
    info = rng.integers(0, 256, (4, K // 8), dtype=np.uint8)
    code = orc.encode(N, fz, info)
    bits = np.unpackbits(code, axis=1, bitorder="big")
    mask = np.ones(N, bool)
    mask[fz] = False
    assert np.array_equal(np.packbits(bits[:, mask], axis=1, bitorder="big"), info)


def test_crc_kats(orc):
    """python/qa_pypolar_detector.py:51-442, README.md:81-91"""
    t = np.frombuffer(b"Test", np.uint8)

    def gen(kind, data, cb):
        full = np.zeros((1, len(data) + cb), np.uint8)
        full[0, :len(data)] = data
        return orc.crc_generate(kind, full)[0, -cb:].tolist()

    assert orc.crc_calculate(8, np.frombuffer(b"TestFooB", np.uint8), 64) == 0xC2
    assert gen(32, t, 4) == [0x8C, 0x2D, 0xE2, 0x19]
    assert gen(16, t, 2) == [0x28, 0x88]
    assert gen(17, t, 2) == [0xAC, 0x48]
    assert gen(24, np.frombuffer(b"123456789", np.uint8), 3) == [0xF4, 0x82, 0x79]


def test_crc_golden(orc):
    data = GOLD["crc_data"]
    for bits, typ, cb in [(6, "crc", 1), (8, "crc", 1), (11, "crc", 2), (16, "crc", 2), (16, "crcnr", 2), (24, "crc", 3),
                          (32, "crc", 4)]:
        kind = orc.crc_kind(bits, typ)
        full = np.zeros((32, 24 + cb), np.uint8)
        full[:, :24] = data
        gen = orc.crc_generate(kind, full)
        assert np.array_equal(gen, GOLD[f"crc_{bits}_{typ}_gen"]), (bits, typ)
        assert orc.crc_check(kind, gen).all()
        gen[:, 2] ^= 1
        assert not orc.crc_check(kind, gen).any()
        if bits != 32:
            got = [orc.crc_calculate(kind, data[i], 35) for i in range(32)]
            assert got == GOLD[f"crc_{bits}_{typ}_calc35"].tolist()


def test_quantizer_kat_and_golden(orc):
    """test/polarcode/bitcontainertest.cpp:73-96: round(clamp(4 i + 0.3))"""
    x = np.array([[4 * i + 0.3 for i in range(-32, 32)]], np.float32)
    assert np.array_equal(orc.quantize(64, x)[0], np.clip(np.rint(x[0]), -128, 127).astype(np.int8))
    assert np.array_equal(orc.quantize(64, GOLD["quant_in"]), GOLD["quant_out_64"])
    assert np.array_equal(orc.quantize(16, GOLD["quant_in"].reshape(-1, 16)), GOLD["quant_out_16"])


# ---- differential against the unmodified reference (skipped where oracle/_ref is absent) -------------------

@pytest.mark.parametrize("N,K", [(1024, 512), (64, 40), (128, 96), (512, 128), (4096, 2048)])
def test_differential_decode(orc, ref, N, K):
    from conftest import awgn_llr_i8
    rng = np.random.default_rng(N)
    fz = ref.frozen_bits(N, K, 0.0)
    assert np.array_equal(fz, orc.frozen_bits(N, K, 0.0))
    B = 400 if N <= 1024 else 60
    llr = np.concatenate([awgn_llr_i8(rng, B, N, 1.0, K / N, 10.0), awgn_llr_i8(rng, B, N, 3.0, K / N, 50.0),
                          rng.choice(np.array([-128, -127, 127, 0], np.int8), (B, N)),
                          rng.integers(-128, 128, (B, N), dtype=np.int8)])
    exp, eok = ref.decode_batch(N, 1, fz, "char", llr)
    got = orc.fastssc_i8_decode(N, fz, llr)
    assert np.array_equal(got, exp)
    assert np.array_equal(orc.crc_check(8, got), eok)


def test_differential_construction(orc, ref):
    """python/qa_pypolar_encoder.py:116-137 shape: several design SNRs, rates and lengths"""
    for N in (64, 256, 1024):
        for K in (N // 8, N // 4, N // 2, 3 * N // 4):
            for d in (-1.5, 0.0, 1.0, 2.5):
                assert np.array_equal(orc.frozen_bits(N, K, d), ref.frozen_bits(N, K, d)), (N, K, d)
            assert np.array_equal(orc.frozen_bits(N, K, 0.0, "BE"), ref.frozen_bits(N, K, 0.0, "BE"))


def test_reference_float_list_decoder_keeps_metric_across_frames(ref):
    """Pins a reference defect the GPU parity tests route around: SclAvxFloat never resets the metric of path 0 between
    decode() calls (PathList::setFirstPath, scl_avx_float.cpp:103-109, assigns no metric; only the constructor zeroes
    it), so every frame starts from the previous frame's best metric. In exact arithmetic (integer-valued LLRs) that
    offset changes nothing; with continuous LLRs the float rounding of every later metric depends on the decoder's
    history and near-tie decisions flip, more and more often as |metric| grows. A fresh object per frame is history
    free -- that is what k_scl_f32 (metric starts at 0 for every frame) is compared with at scale (tools/parity_sweep.py)."""
    N, K, L = 256, 128, 4
    fz = ref.frozen_bits(N, K, 0.0)
    rng = np.random.default_rng(0)
    llr = (4.0 * (1 + rng.standard_normal((12000, N)))).astype(np.float32)
    long_lived, _ = ref.decode_batch(N, L, fz, "float", llr, crc_bits=0, threads=1)
    fresh, _ = ref.decode_batch(N, L, fz, "float", llr, crc_bits=0, threads=4, fresh=True)
    bad = np.flatnonzero((long_lived != fresh).any(axis=1))
    assert 0 < len(bad) < 0.05 * len(llr)
    assert (bad >= len(llr) // 2).sum() > (bad < len(llr) // 2).sum()  # drifts: worse the longer the object lives
    ints = np.rint(llr).astype(np.float32)
    a, _ = ref.decode_batch(N, L, fz, "float", ints, crc_bits=0, threads=1)
    b, _ = ref.decode_batch(N, L, fz, "float", ints, crc_bits=0, threads=4, fresh=True)
    assert np.array_equal(a, b)
    # the int8 list decoder has the same carry-over but integer metrics: no effect
    i8 = np.clip(np.rint(llr * 2.5), -128, 127).astype(np.int8)
    a, _ = ref.decode_batch(N, L, fz, "char", i8, crc_bits=0, threads=1)
    b, _ = ref.decode_batch(N, L, fz, "char", i8, crc_bits=0, threads=4, fresh=True)
    assert np.array_equal(a, b)


# ---- list decoder restatement (oracle/pc_oracle_scl.c) -------------------------------------------------------------
GOLD_LIST = np.load(os.path.join(os.path.dirname(__file__), "golden", "reference_vectors_list.npz"))
SCLC_KEYS = sorted(k[:-4] for k in GOLD_LIST.files if k.startswith("sclc_") and k.endswith("_llr"))


def _crc_kind(bits, typ):
    return 17 if (bits == 16 and typ == "crcnr") else bits


@pytest.mark.parametrize("key", SCLC_KEYS)
def test_scl_restatement_golden(orc, key):
    """SclFipChar outputs recorded by tests/golden/make_golden_list.py"""
    _, N, K, L, cb, ct = key.split("_")
    N, L = int(N), int(L)
    got, ok = orc.scl_i8_decode(N, GOLD_LIST[key + "_frozen"], L, GOLD_LIST[key + "_llr"], crc=_crc_kind(int(cb), ct))
    assert np.array_equal(got, GOLD_LIST[key + "_info"])
    assert np.array_equal(ok, GOLD_LIST[key + "_ok"])


@pytest.mark.parametrize("N,K,L", [(1024, 512, 8), (256, 128, 4), (64, 32, 2), (128, 96, 16), (512, 128, 32), (16, 8, 4)])
def test_scl_restatement_differential(orc, ref, N, K, L):
    """random noisy all-zero codewords, saturated and uniformly random LLRs: the adversarial inputs for the partial
    sort and the weak-LLR search (many equal metrics)"""
    from conftest import awgn_llr_i8
    rng = np.random.default_rng(N + L)
    fz = ref.frozen_bits(N, K, 0.0)
    B = 60 if N <= 256 else 24
    llr = np.concatenate([awgn_llr_i8(rng, B, N, 1.0, K / N, 10.0), awgn_llr_i8(rng, B, N, 3.0, K / N, 50.0),
                          rng.choice(np.array([-128, -127, 127, 0], np.int8), (B // 2, N)),
                          rng.integers(-128, 128, (B // 2, N), dtype=np.int8)])
    exp, eok = ref.decode_batch(N, L, fz, "char", llr, crc_bits=8)
    got, ok = orc.scl_i8_decode(N, fz, L, llr, crc=8)
    assert np.array_equal(got, exp) and np.array_equal(ok, eok)
