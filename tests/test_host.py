"""CPU tests of the host side: the C-ABI library loads and exports what include/polar_b200.h declares, the host-only
entry points work without a GPU, compute entry points fail loudly without one, and the device decoder core (compiled
for the host by tests/host_emul.cpp -- a debug harness, not a product path) agrees with the oracle."""
import ctypes
import os
import re
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLD = np.load(os.path.join(os.path.dirname(__file__), "golden", "reference_vectors.npz"))


def _has_gpu():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


def test_header_symbols_exported(pcb):
    hdr = open(os.path.join(ROOT, "include", "polar_b200.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    declared = set(re.findall(r"\b(pc_[A-Za-z0-9_]+)\s*\(", hdr))
    assert len(declared) >= 30
    lib = pcb.capi.lib()
    for name in sorted(declared):
        assert hasattr(lib, name), f"{name} declared in include/polar_b200.h but not exported"
    assert declared == set(pcb.capi.SIGNATURES), declared ^ set(pcb.capi.SIGNATURES)
    assert b"sm_100a" in lib.pc_version()


def test_frozen_bits_host_entry(pcb, orc):
    for key in GOLD.files:
        if key.startswith("frozen_"):
            _, t, N, K, d = key.split("_")
            assert pcb.frozen_bits(int(N), int(K), float(d), t) == GOLD[key].tolist(), key
    with pytest.raises(ValueError):
        pcb.frozen_bits(100, 50, 0.0)  # constructor.cpp:25-31: std::invalid_argument for a non power of two


def test_detector_factory_rules(pcb):
    """ErrorDetection::create (errordetector.cpp:27-88)"""
    L = pcb.capi.lib()
    assert L.pc_crc_kind(16, b"crcnr") == 17 and L.pc_crc_kind(16, b"CRC") == 16 and L.pc_crc_kind(0, b"crc") == 0
    assert L.pc_crc_kind(7, b"crc") < 0
    assert [L.pc_crc_bytes(k) for k in (0, 6, 8, 11, 16, 17, 24, 32)] == [0, 1, 1, 2, 2, 2, 3, 4]
    d = pcb.Detector(11, "crc")
    assert d.getCheckBitCount() == 11 and d.getType() == "CRCNR"  # crc11nr.h:36
    # getType() / getCheckBitCount() of every detector the factory builds (values printed by the reference's pybind module)
    for size, typ, name in ((0, "crc", "DUMMY"), (6, "crc", "CRCNR"), (8, "crc", "CRC"), (16, "crc", "CRC"),
                            (16, "crcnr", "CRCNR"), (24, "crc", "CRCNR"), (32, "crc", "CRC")):
        d = pcb.Detector(size, typ)
        assert (d.getType(), d.getCheckBitCount()) == (name, size)


def test_decoder_kind_mapping(pcb):
    """Decoding::create / makeDecoder (decoder.cpp:26-87)"""
    from polar_codes_b200.pypolar import _kind_from_string as k
    c = pcb.capi
    assert k("char", 1) == c.PC_DEC_FASTSSC_I8 and k("CHAR", 8) == c.PC_DEC_SCL_I8
    assert k("float", 1) == c.PC_DEC_FASTSSC_F32 and k("float", 4) == c.PC_DEC_SCL_F32
    assert k("mixed", 4) == c.PC_DEC_ADAPTIVE_F32 and k("mixed", 1) == c.PC_DEC_FASTSSC_F32
    with pytest.raises(RuntimeError):
        k("bogus", 1)


@pytest.mark.skipif(_has_gpu(), reason="checks the no-GPU failure mode")
def test_no_cpu_fallback(pcb):
    """the product path must fail loudly when no CUDA device is usable"""
    with pytest.raises(pcb.capi.PolarError):
        pcb.PolarDecoder(8, 1, [0, 1, 2, 4], "char")
    assert "no usable CUDA device" in pcb.capi.last_error()


@pytest.fixture(scope="module")
def emul():
    build = os.path.join(ROOT, "tests", "_build")
    os.makedirs(build, exist_ok=True)
    so = os.path.join(build, "libhost_emul.so")
    srcs = [os.path.join(ROOT, "tests", "host_emul.cpp"), os.path.join(ROOT, "polar-codes_b200", "csrc", "pc_program.cpp")]
    deps = srcs + [os.path.join(ROOT, "polar-codes_b200", "csrc", f) for f in
                   ("pc_fastssc_i8_core.h", "pc_fastssc_f32_core.h", "pc_swar.h", "pc_crc.h", "pc_program.h")]
    if not os.path.exists(so) or os.path.getmtime(so) < max(os.path.getmtime(d) for d in deps):
        subprocess.check_call(["g++", "-std=c++17", "-O2", "-fPIC", "-shared", "-ffp-contract=off", "-o", so] + srcs)
    return ctypes.CDLL(so)


def _emul_decode(E, N, fz, llr, crc=8, syst=1, mode=0):
    fz = np.ascontiguousarray(fz, dtype=np.uint32)
    llr = np.ascontiguousarray(llr, dtype=np.int8)
    B, K = llr.shape[0], N - len(fz)
    out = np.zeros((B, (K + 7) // 8), np.uint8)
    ok = np.zeros(B, np.uint8)
    vp = ctypes.c_void_p
    r = E.emul_fastssc_i8_decode_mode(N, fz.ctypes.data_as(vp), len(fz), syst, crc, llr.ctypes.data_as(vp),
                                      ctypes.c_long(B), out.ctypes.data_as(vp), ok.ctypes.data_as(vp), mode)
    assert r == 0
    return out, ok


@pytest.mark.parametrize("mode", [0, 1, 2, 3])
def test_device_core_on_host_golden(emul, mode):
    """node program + SWAR arithmetic + placement (shared / tensor memory / arena emulation) against the reference's outputs"""
    for N, K in [(1024, 512), (64, 40), (8, 4), (256, 64), (2048, 1800), (16384, 8192)]:
        out, ok = _emul_decode(emul, N, GOLD[f"dec_{N}_{K}_frozen"], GOLD[f"dec_{N}_{K}_llr"], mode=mode)
        assert np.array_equal(out, GOLD[f"dec_{N}_{K}_info"]), (N, K, mode)
        assert np.array_equal(ok, GOLD[f"dec_{N}_{K}_ok"])
    out, ok = _emul_decode(emul, 1024, GOLD["dec_1024_512_frozen"], GOLD["decns_1024_512_llr"], syst=0, mode=mode)
    assert np.array_equal(out, GOLD["decns_1024_512_info"]) and np.array_equal(ok, GOLD["decns_1024_512_ok"])


@pytest.mark.parametrize("N,K", [(1024, 512), (32, 16), (16, 8), (128, 100), (1024, 1), (1024, 1023), (4096, 1024)])
def test_device_core_on_host_vs_oracle(emul, orc, N, K):
    rng = np.random.default_rng(N + K)
    fz = orc.frozen_bits(N, K, 0.0)
    llr = np.concatenate([np.clip(np.rint(rng.normal(1, 1, (150, N)) * a), -128, 127).astype(np.int8) for a in (3, 10, 60)] +
                         [rng.choice(np.array([-128, 127, 0, -1], np.int8), (100, N))])
    exp = orc.fastssc_i8_decode(N, fz, llr)
    for mode in (0, 1):
        out, ok = _emul_decode(emul, N, fz, llr, mode=mode)
        assert np.array_equal(out, exp)
        assert np.array_equal(ok, orc.crc_check(8, exp))


def test_swar_primitives_exhaustive(orc):
    """pc_swar.h f / g / adds on all 65536 byte pairs, host build of the same header (tools-free: via the emulated decoder
    of a length-2 ... covered by the decode tests above); here the algebra used by sw_f is restated in numpy"""
    a = np.repeat(np.arange(-128, 128), 256)
    b = np.tile(np.arange(-128, 128), 256)
    x = np.abs(a)
    x = np.where(x == 128, 127, x)
    y = np.abs(b)
    m = (x + y - np.abs(x - y)) >> 1
    mag = (m + 1 + np.abs(m - 1)) >> 1
    out = np.where((a < 0) ^ (b < 0), -mag, mag)
    assert np.array_equal(out.astype(np.int8), orc.f_i8(a.astype(np.int8), b.astype(np.int8)))
    assert (x + y).max() <= 255 and (m + 1 + np.abs(m - 1)).max() <= 254


@pytest.mark.parametrize("N,K", [(8, 4), (16, 8), (32, 16), (64, 40), (128, 100), (256, 128), (512, 256), (1024, 512), (1024, 900),
                                 (1024, 100), (2048, 1024)])
def test_float_soft_container_on_host_vs_reference(emul, ref, N, K):
    """the fp32 core with its soft-output context (FSoftCtx, what k_fastssc_f32_soft runs per thread) on the CPU against
    FloatContainer::getSoftBits of the unmodified reference, compared as 32-bit patterns (inner words are XORs of floats)"""
    rng = np.random.default_rng(7 * N + K)
    for dsnr in (0.0, 2.0):
        fz = np.ascontiguousarray(ref.frozen_bits(N, K, dsnr), dtype=np.uint32)
        llr = np.concatenate([(a * (1.0 + s * rng.standard_normal((40, N)))).astype(np.float32) for a, s in ((2.0, 1.0), (8.0, 0.7), (1.0, 3.0))] +
                             [rng.choice(np.array([0.0, -0.0, 1.0, -1.0, 2.5], np.float32), (40, N))])
        B = llr.shape[0]
        got = np.zeros((B, N), np.uint32)
        vp = ctypes.c_void_p
        r = emul.emul_fastssc_f32_soft(N, fz.ctypes.data_as(vp), len(fz), llr.ctypes.data_as(vp), ctypes.c_long(B), got.ctypes.data_as(vp))
        if r == 3:
            with pytest.raises(RuntimeError):
                ref.decode_soft_f32(N, fz, llr)
            continue
        assert r == 0
        esc, _ = ref.decode_soft_f32(N, fz, llr)
        exp = esc.view(np.uint32)
        bad = (got != exp).any(axis=1)
        assert not bad.any(), f"{bad.sum()} of {B} frames differ, first: frame {np.argmax(bad)} at {np.nonzero(got[np.argmax(bad)] != exp[np.argmax(bad)])[0][:8]}"
