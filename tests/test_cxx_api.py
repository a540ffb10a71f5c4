"""The C++ class mirror include/polarcode_b200.hpp (PolarCode::Encoding::ButterflyFipPacked, Decoding::create /
Decoder, ErrorDetection::create / Detector, Construction::frozen_bits) driven through tests/cxx_api.cpp the way a
caller of the reference's classes would, and compared with the unmodified reference (oracle/_ref)."""
import ctypes
import os
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
VP = ctypes.c_void_p


@pytest.fixture(scope="module")
def cxx(pcb):
    pcb.capi.lib()
    build = os.path.join(ROOT, "tests", "_build")
    os.makedirs(build, exist_ok=True)
    so = os.path.join(build, "libcxx_api.so")
    deps = [os.path.join(ROOT, "tests", "cxx_api.cpp"), os.path.join(ROOT, "include", "polarcode_b200.hpp"),
            os.path.join(ROOT, "include", "polar_b200.h")]
    if not os.path.exists(so) or os.path.getmtime(so) < max(os.path.getmtime(d) for d in deps):
        # -Bsymbolic + hidden visibility: the mirror has the reference's class names ON PURPOSE, and the test process
        # also holds the real libPolarCode.so (RTLD_GLOBAL, the oracle); without this the dynamic linker would bind the
        # harness's inline PolarCode::ErrorDetection::create(...) to the reference's symbol of the same name
        subprocess.check_call(["g++", "-std=c++17", "-O2", "-fPIC", "-shared", "-Wall", "-fvisibility-inlines-hidden",
                               "-Wl,-Bsymbolic", "-o", so, deps[0],
                               "-L" + os.path.join(ROOT, "polar-codes_b200"), "-lpolar_b200",
                               "-Wl,-rpath,$ORIGIN/../../polar-codes_b200"])
    return ctypes.CDLL(so)


def _p(a):
    return a.ctypes.data_as(VP)


def test_cxx_frozen_bits_and_errors(cxx, ref):
    out = np.zeros(4096, np.uint32)
    n = ctypes.c_int(0)
    for N, K, d, t in ((1024, 512, 0.0, "BB"), (64, 40, 1.0, "BB"), (512, 128, 0.0, "BE")):
        assert cxx.cxx_frozen_bits(N, K, ctypes.c_float(d), t.encode(), _p(out), ctypes.byref(n)) == 0
        assert out[:n.value].tolist() == list(ref.frozen_bits(N, K, d, t))
    # constructor.cpp:25-31: std::invalid_argument for a block length that is not a power of two
    assert cxx.cxx_frozen_bits(100, 50, ctypes.c_float(0.0), b"BB", _p(out), ctypes.byref(n)) == 1


@pytest.mark.gpu
@pytest.mark.parametrize("size,typ", [(0, "crc"), (6, "crc"), (8, "crc"), (11, "crc"), (16, "crc"), (16, "crcnr"),
                                      (24, "crc"), (32, "crc")])
def test_cxx_detector_matches_reference(cxx, ref, size, typ):
    rng = np.random.default_rng(size)
    for nbytes in (8, 16, 64):
        data = rng.integers(0, 256, nbytes + 8, dtype=np.uint8)
        buf = data.copy()
        okf, bad, calc, bits = ctypes.c_int(0), ctypes.c_int(0), ctypes.c_ulonglong(0), ctypes.c_int(0)
        mode = ctypes.create_string_buffer(32)
        assert cxx.cxx_detector(size, typ.encode(), _p(buf), nbytes, ctypes.byref(okf), ctypes.byref(bad),
                                ctypes.byref(calc), mode, ctypes.byref(bits)) == 0
        cb = (size + 7) // 8
        exp = ref.crc_generate(size, typ, data[None, :nbytes + cb].copy())[0]
        assert np.array_equal(buf[:nbytes + cb], exp)
        # canary: generate(data, total) replaces the LAST bytes inside the block and touches nothing behind it
        assert np.array_equal(buf[nbytes + cb:], data[nbytes + cb:])
        assert okf.value == 1 and (bad.value == 0 or size == 0)
        assert bits.value == size
    # errordetector.cpp:41: std::logic_error("CRC INVALID SIZE!")
    assert cxx.cxx_detector(7, b"crc", _p(buf), 8, ctypes.byref(okf), ctypes.byref(bad), ctypes.byref(calc), mode,
                            ctypes.byref(bits)) == 2


def test_cxx_no_gpu_is_an_error(cxx):
    """no CPU fallback: constructing a decoder without a usable device throws std::runtime_error"""
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    fz = np.array([0, 1, 2, 4], np.uint32)
    out, ok = np.zeros(8, np.uint8), np.zeros(1, np.uint8)
    mode, ls = ctypes.create_string_buffer(32), ctypes.c_int(0)
    llr = np.zeros(8, np.int8)
    assert cxx.cxx_decode(8, 1, _p(fz), 4, b"char", -1, b"crc", 1, 0, _p(llr), ctypes.c_long(1), _p(out), _p(ok), 0, mode,
                          ctypes.byref(ls)) == 3


@pytest.mark.gpu
@pytest.mark.parametrize("N,K,crc,typ,syst", [(1024, 512, 0, "crc", 1), (1024, 512, 16, "crcnr", 1), (64, 40, 8, "crc", 1),
                                              (256, 128, 24, "crc", 0), (2048, 1024, 32, "crc", 1)])
def test_cxx_encoder_matches_reference(cxx, ref, N, K, crc, typ, syst):
    rng = np.random.default_rng(N + K + crc)
    fz = np.ascontiguousarray(ref.frozen_bits(N, K, 0.0), dtype=np.uint32)
    B = 40
    info = rng.integers(0, 256, (B, K // 8), dtype=np.uint8)
    exp_code, exp_info = ref.encode_batch(N, fz, info, crc_bits=crc, crc_type=typ, systematic=bool(syst))
    for batch in (0, 1):
        mine, code = info.copy(), np.zeros((B, N // 8), np.uint8)
        mode = ctypes.create_string_buffer(32)
        assert cxx.cxx_encode(N, _p(fz), len(fz), crc, typ.encode(), syst, _p(mine), _p(code), ctypes.c_long(B), batch, mode) == 0
        assert np.array_equal(code, exp_code) and np.array_equal(mine, exp_info)


@pytest.mark.gpu
@pytest.mark.parametrize("N,K,L,kind,crc,typ", [(1024, 512, 1, "char", -1, "crc"), (1024, 512, 1, "float", 8, "crc"),
                                                (64, 40, 4, "mixed", -1, "crc"), (512, 128, 8, "float", 24, "crc"),
                                                (256, 128, 4, "CHAR", 16, "crc")])
def test_cxx_decoder_matches_reference(cxx, ref, N, K, L, kind, crc, typ):
    from conftest import awgn_llr_i8
    rng = np.random.default_rng(N + K + L)
    fz = np.ascontiguousarray(ref.frozen_bits(N, K, 0.0), dtype=np.uint32)
    B = 48
    llr8 = awgn_llr_i8(rng, B, N, 1.5, K / N)
    is_float = "char" not in kind.lower()
    llr = llr8.astype(np.float32) * np.float32(0.37) if is_float else llr8
    exp, eok = ref.decode_batch(N, L, fz, kind, llr, crc_bits=crc, crc_type=typ)
    for batch in (0, 1):
        out, ok = np.zeros((B, K // 8), np.uint8), np.zeros(B, np.uint8)
        mode, ls = ctypes.create_string_buffer(32), ctypes.c_int(0)
        assert cxx.cxx_decode(N, L, _p(fz), len(fz), kind.encode(), crc, typ.encode(), 1, int(is_float), _p(llr),
                              ctypes.c_long(B), _p(out), _p(ok), batch, mode, ctypes.byref(ls)) == 0
        assert np.array_equal(out, exp) and np.array_equal(ok, eok)
        assert ls.value == L
        # makeDecoder installs CRC-8 (decoder.cpp:85)
        assert mode.value.decode() == ("CRC-8" if crc < 0 else {8: "CRC-8", 24: "CRCNR-24", 16: "CRC-16"}[crc])
    # decoder.cpp:45: std::logic_error("Unknown PolarDecoder type!")
    assert cxx.cxx_decode(N, L, _p(fz), len(fz), b"bogus", crc, typ.encode(), 1, 0, _p(llr8), ctypes.c_long(1), _p(out),
                          _p(ok), 0, mode, ctypes.byref(ls)) == 2


@pytest.mark.gpu
def test_cxx_puncturer_matches_reference(cxx, ref):
    N, M = 256, 200
    rng = np.random.default_rng(1)
    f = np.ascontiguousarray(ref.frozen_bits(N, 128, 0.0), np.uint32)
    pos = np.zeros(M, np.uint32)
    parent = ctypes.c_long(0)
    x = rng.standard_normal(N).astype(np.float32)
    bits = rng.integers(0, 256, N // 8, dtype=np.uint8)
    punct, packed, dep = np.zeros(M, np.float32), np.zeros(M // 8, np.uint8), np.zeros(N, np.float32)
    cxx.cxx_puncturer.argtypes = [ctypes.c_long, VP, ctypes.c_int, VP, ctypes.POINTER(ctypes.c_long), VP, VP, VP, VP, VP]
    assert cxx.cxx_puncturer(M, _p(f), len(f), _p(pos), ctypes.byref(parent), _p(x), _p(punct), _p(bits), _p(packed), _p(dep)) == 0
    assert parent.value == N and pos.tolist() == ref.puncturer(M, f, "positions").tolist()
    assert np.array_equal(punct, ref.puncturer(M, f, "puncture", x[None])[0])
    assert np.array_equal(packed, ref.puncturer(M, f, "packed", bits[None])[0])
    assert np.array_equal(dep, ref.puncturer(M, f, "depuncture", punct[None])[0])
    assert cxx.cxx_puncturer(100, _p(f), 10, _p(pos), ctypes.byref(parent), _p(x), _p(punct), _p(bits), _p(packed), _p(dep)) == 5


@pytest.mark.gpu
def test_cxx_soft_getters_match_reference(cxx, ref):
    """getSoftCodeword / getSoftInformation of the class mirror: int8 container for 'char', float patterns for 'float';
    the list decoders throw std::logic_error"""
    N, K = 256, 128
    fz = np.ascontiguousarray(ref.frozen_bits(N, K, 0.0), dtype=np.uint32)
    rng = np.random.default_rng(9)
    y = (3.0 * (1.0 + 0.8 * rng.standard_normal((24, N)))).astype(np.float32)
    sc, si = np.zeros((24, N), np.float32), np.zeros((24, K), np.float32)
    assert cxx.cxx_soft(N, 1, _p(fz), len(fz), b"float", 1, _p(y), ctypes.c_long(24), _p(sc), _p(si), 4) == 0
    esc, esi = ref.decode_soft_f32(N, fz, y)
    assert np.array_equal(sc.view(np.uint32), esc.view(np.uint32)) and np.array_equal(si.view(np.uint32), esi.view(np.uint32))
    q = np.clip(np.rint(y * 4), -128, 127).astype(np.int8)
    bc, bi = np.zeros((24, N), np.int8), np.zeros((24, K), np.int8)
    assert cxx.cxx_soft(N, 1, _p(fz), len(fz), b"char", 0, _p(q), ctypes.c_long(24), _p(bc), _p(bi), 1) == 0
    ebc, ebi = ref.decode_soft(N, fz, q)
    assert np.array_equal(bc, ebc) and np.array_equal(bi, ebi)
    assert cxx.cxx_soft(N, 4, _p(fz), len(fz), b"char", 0, _p(q), ctypes.c_long(1), _p(bc), _p(bi), 1) == 2


@pytest.mark.gpu
@pytest.mark.parametrize("N,K", [(256, 128), (1024, 512), (2048, 1500)])
def test_cxx_encoder_codeword_hooks_match_reference(cxx, ref, N, K):
    """Encoder::setCodeword / setCharCodeword / setFloatCodeword / clearFrozenBits (encoder.cpp:56-77) + encode():
    the transform of a ready codeword, both modes, against the unmodified reference (N >= 256: below that the
    reference's insertCharBits misplaces the bits in its 256-bit vector)"""
    import refshim
    L = refshim.lib()
    L.ref_encode_codeword.restype = ctypes.c_int
    L.ref_encode_codeword.argtypes = [ctypes.c_int, VP, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_int, VP, VP]
    fz = np.ascontiguousarray(ref.frozen_bits(N, K, 0.0), dtype=np.uint32)
    rng = np.random.default_rng(N)
    bits = rng.integers(0, 2, N, dtype=np.uint8)
    packed = np.packbits(bits, bitorder="big")
    chars = np.where(bits == 1, rng.integers(-128, 0, N), rng.integers(0, 128, N)).astype(np.int8)
    floats = np.where(bits == 1, -1.0, 1.0).astype(np.float32) * rng.uniform(0.1, 5.0, N).astype(np.float32)
    for syst in (0, 1):
        for clear in (0, 1):
            for form, arr in ((0, packed), (1, chars), (2, floats)):
                got, exp = np.zeros(N // 8, np.uint8), np.zeros(N // 8, np.uint8)
                a1, a2 = arr.copy(), arr.copy()
                assert cxx.cxx_encode_codeword(N, _p(fz), len(fz), syst, form, clear, _p(a1), _p(got)) == 0
                assert L.ref_encode_codeword(N, _p(fz), len(fz), syst, form, clear, _p(a2), _p(exp)) == 0
                assert np.array_equal(got, exp), (N, K, syst, clear, form)
