"""GPU parity: encoder, detectors, quantiser, channel generator and error counting through the C-ABI,
against the unmodified reference (oracle/_ref) and our C restatement (oracle/pc_oracle.c). Bit-exact."""
import ctypes

import numpy as np
import pytest

from conftest import awgn_llr_i8

pytestmark = pytest.mark.gpu

CRCS = [(0, "crc"), (6, "crc"), (8, "crc"), (11, "crc"), (16, "crc"), (16, "crcnr"), (24, "crc"), (32, "crc")]


@pytest.mark.parametrize("N,K", [(1024, 512), (8, 4), (16, 8), (32, 16), (64, 40), (128, 32), (256, 128), (512, 128),
                                 (2048, 1024), (4096, 3072), (16384, 8192), (65536, 32768), (1024, 1016)])
@pytest.mark.parametrize("systematic", [True, False])
def test_encoder_parity(pcb, ref, orc, N, K, systematic):
    if N == 65536:
        fz = orc.frozen_bits(N, K, 0.0)
    else:
        fz = ref.frozen_bits(N, K, 0.0)
    rng = np.random.default_rng(N + K)
    B = 777 if N <= 4096 else 37
    info = rng.integers(0, 256, (B, (K + 7) // 8), dtype=np.uint8)
    info[0] = 0
    info[1] = 0xFF
    if K % 8:
        info[:, -1] &= (0xFF00 >> (K % 8)) & 0xFF
    enc = pcb.PolarEncoder(N, list(fz))
    enc.setSystematic(systematic)
    got = enc.encode_batch(info)
    exp = orc.encode(N, fz, info, systematic=systematic)
    assert np.array_equal(got, exp)
    if N >= 256 and N <= 16384:  # the reference packs N < 256 into one 256-bit vector; same bits, checked via the C restatement
        exp2, _ = ref.encode_batch(N, fz, info, crc_bits=0, systematic=systematic)
        assert np.array_equal(got, exp2)


@pytest.mark.parametrize("crc_bits,crc_type", CRCS)
def test_encoder_with_detector(pcb, ref, crc_bits, crc_type):
    N, K = 1024, 512
    fz = ref.frozen_bits(N, K, 0.0)
    rng = np.random.default_rng(crc_bits)
    info = rng.integers(0, 256, (500, K // 8), dtype=np.uint8)
    enc = pcb.PolarEncoder(N, list(fz))
    enc.setErrorDetection(crc_bits, crc_type)
    before = info.copy()
    code, icrc = enc.encode_batch(info, return_info=True)
    exp_code, exp_info = ref.encode_batch(N, fz, info, crc_bits=crc_bits, crc_type=crc_type)
    assert np.array_equal(info, before)  # the caller's buffer is not written (the reference writes the CRC in place)
    assert np.array_equal(icrc, exp_info)
    assert np.array_equal(code, exp_code)


def test_encoder_kats(pcb):
    """reference test/polarcode/encodingtest.cpp:26-90"""
    for n in range(3, 11):
        N = 1 << n
        enc = pcb.PolarEncoder(N, [])  # K = N: u = e_{N-1} -> all-ones codeword (non-systematic)
        enc.setSystematic(False)
        u = np.zeros(N // 8, np.uint8)
        u[-1] = 1
        assert (enc.encode_vector(u) == 0xFF).all()
    fz = pcb.frozen_bits(16, 8, 0.0)
    enc = pcb.PolarEncoder(16, fz)
    assert enc.encode_vector(np.array([0xFF], np.uint8)).tolist() == [0xFF, 0xFF]


@pytest.mark.parametrize("crc_bits,crc_type", CRCS)
@pytest.mark.parametrize("nbytes", [8, 9, 64, 131])
def test_detector_batch(pcb, ref, orc, crc_bits, crc_type, nbytes):
    if crc_bits == 32 and nbytes % 4:
        pytest.skip("CRC32 of the reference needs whole words")
    rng = np.random.default_rng(nbytes)
    det = pcb.Detector(crc_bits, crc_type)
    data = rng.integers(0, 256, (300, nbytes), dtype=np.uint8)
    gen = det.generate_batch(data)
    cb = gen.shape[1] - nbytes
    full = np.zeros((300, nbytes + cb), np.uint8)
    full[:, :nbytes] = data
    kind = orc.crc_kind(crc_bits, crc_type)
    exp = orc.crc_generate(kind, full) if crc_bits else full
    assert np.array_equal(gen, exp)
    if crc_bits:
        assert np.array_equal(gen, ref.crc_generate(crc_bits, crc_type, full))
    assert det.check_batch(gen).all()
    if crc_bits:
        bad = gen.copy()
        bad[:, 3] ^= 0x10
        got = det.check_batch(bad)
        assert np.array_equal(got.astype(np.uint8), orc.crc_check(kind, bad))
        assert not got.any()


def test_detector_kats(pcb):
    """reference python/qa_pypolar_detector.py:51-442, README.md:81-91"""
    D = pcb.Detector
    t = np.frombuffer(b"Test", np.uint8)
    assert D(8, "crc").calculate(np.frombuffer(b"TestFooB", np.uint8), 64) == 0xC2
    assert D(8, "crc").generate(np.frombuffer(b"TestFooB", np.uint8))[-1] == 0xC2
    assert D(32, "crc").generate(t).tolist()[-4:] == [0x8C, 0x2D, 0xE2, 0x19]
    assert D(16, "crc").generate(t).tolist()[-2:] == [0x28, 0x88]
    assert D(16, "crcnr").generate(t).tolist()[-2:] == [0xAC, 0x48]
    assert D(24, "crc").generate(np.frombuffer(b"123456789", np.uint8)).tolist()[-3:] == [0xF4, 0x82, 0x79]
    assert D(0, "crc").check(t)
    with pytest.raises(ValueError):
        D(7, "crc")
    # bit-granular calculate(): the 35-bit and 13-bit messages of qa_pypolar_detector.py:310-319,367-376,433-442
    b35 = np.packbits(np.array([0, 0, 1, 0, 0, 0, 1, 0, 0, 0, 1, 0, 0, 1, 1, 1, 0, 0, 0, 1, 1, 1, 1, 0, 1, 1, 1, 1, 0, 0, 1,
                                0, 0, 1, 0], dtype=np.uint8))
    b13 = np.packbits(np.array([1, 0, 1, 1, 1, 1, 0, 1, 0, 0, 0, 1, 0], dtype=np.uint8))
    assert D(24, "crcnr").calculate(b35, 35) == 0x1FB065 and D(24, "crcnr").calculate(b13, 13) == 0xF70030
    assert D(6, "crcnr").calculate(b35, 35) == 0x02 and D(6, "crcnr").calculate(b13, 13) == 0x2F
    assert D(11, "crcnr").calculate(b35, 35) == 0x0608 and D(11, "crcnr").calculate(b13, 13) == 0x06C8
    # generate(): CRC-6 / CRC-11 occupy whole bytes (:334-365, :392-431)
    m1, m2 = np.frombuffer(b"123456789", np.uint8), np.frombuffer(b"CommsIsAComplicatedMatter", np.uint8)
    assert D(6, "crcnr").generate(m1).tolist()[-1:] == [0x15] and D(6, "crcnr").generate(m2).tolist()[-1:] == [0x04]
    assert D(11, "crcnr").generate(m1).tolist()[-2:] == [0x05, 0xCA] and D(11, "crcnr").generate(m2).tolist()[-2:] == [0x04, 0x4E]


@pytest.mark.parametrize("crc_bits,crc_type", CRCS[1:])
def test_detector_calculate_bits(pcb, orc, crc_bits, crc_type):
    rng = np.random.default_rng(crc_bits + 100)
    det = pcb.Detector(crc_bits, crc_type)
    kind = orc.crc_kind(crc_bits, crc_type)
    for bits in (8, 13, 35, 64, 100, 257):
        data = rng.integers(0, 256, 160, dtype=np.uint8)
        if crc_bits == 32:
            bits = (bits // 8) * 8
        assert det.calculate(data, bits) == orc.crc_calculate(kind, data, bits)


@pytest.mark.parametrize("N", [4, 8, 16, 32, 1024])
def test_quantizer(pcb, ref, orc, N):
    """reference test/polarcode/bitcontainertest.cpp:73-96 plus the edge values of each rounding path"""
    import torch
    rng = np.random.default_rng(N)
    x = (rng.standard_normal((64, N)) * 60).astype(np.float32)
    x[0, :] = np.resize(np.array([0.5, 1.5, 2.5, -0.5, -1.5, -2.5, 127.5, -128.5, 1e10, -1e10, 3e9, np.nan, np.inf,
                                  -np.inf, 126.5, -127.5], np.float32), N)
    x[1, :] = np.resize(np.array([4 * i + 0.3 for i in range(-32, 32)], np.float32), N)
    d_in = torch.from_numpy(x).cuda()
    d_out = torch.empty((64, N), dtype=torch.int8, device="cuda")
    pcb.capi.check(pcb.capi.lib().pc_quantize_batch(N, pcb.capi.ptr(d_in), pcb.capi.ptr(d_out), 64, None))
    got = d_out.cpu().numpy()
    assert np.array_equal(got, orc.quantize(N, x))
    if N >= 8:
        assert np.array_equal(got, ref.quantize(N, x))


def test_float_input_to_char_decoder(pcb, ref):
    N, K = 1024, 512
    fz = ref.frozen_bits(N, K, 0.0)
    rng = np.random.default_rng(4)
    x = (10.0 * (1.0 + 0.7 * rng.standard_normal((3000, N)))).astype(np.float32)
    dec = pcb.PolarDecoder(N, 1, list(fz), "char")
    got, ok = dec.decode_batch(x, return_ok=True)
    exp, eok = ref.decode_batch(N, 1, fz, "char", x)
    assert np.array_equal(got, exp) and np.array_equal(ok, eok)


def test_channel_generator_and_counting(pcb, ref):
    """device sweep step: generate -> decode -> count; the generated LLRs are replayed through the reference"""
    import torch
    N, K = 1024, 512
    fz = ref.frozen_bits(N, K, 0.0)
    dec = pcb.PolarDecoder(N, 1, list(fz), "char")
    L = pcb.capi.lib()
    B = 20000
    llr = torch.empty((B, N), dtype=torch.int8, device="cuda")
    info = torch.empty((B, K // 8), dtype=torch.uint8, device="cuda")
    pcb.capi.check(L.pc_awgn_generate(dec._h, 0x5EED, 1000, 2.0, 10.0, 0, pcb.capi.ptr(llr), pcb.capi.ptr(info), B, None))
    # shard invariance: the second half generated alone equals the second half of the whole
    llr2 = torch.empty((B // 2, N), dtype=torch.int8, device="cuda")
    info2 = torch.empty((B // 2, K // 8), dtype=torch.uint8, device="cuda")
    pcb.capi.check(L.pc_awgn_generate(dec._h, 0x5EED, 1000 + B // 2, 2.0, 10.0, 0, pcb.capi.ptr(llr2), pcb.capi.ptr(info2),
                                      B // 2, None))
    assert torch.equal(llr[B // 2:], llr2) and torch.equal(info[B // 2:], info2)
    h_info = info.cpu().numpy()
    assert ref.crc_check(8, "crc", h_info).all()  # CRC-8 (the decoder's default detector) is in place
    # noise statistics: LLR = 10 * (x + sigma n), sigma^2 = 1 / (2 * 0.5 * 10^0.2)
    code, _ = ref.encode_batch(N, fz, h_info, crc_bits=0)
    x = 1.0 - 2.0 * np.unpackbits(code[:2000], axis=1, bitorder="big").astype(np.float64)
    noise = llr[:2000].cpu().numpy().astype(np.float64) / 10.0 - x
    sigma = 1.0 / np.sqrt(2 * 0.5 * 10 ** 0.2)
    inner = np.abs(llr[:2000].cpu().numpy().astype(np.int32)) < 120
    assert abs(noise[inner].std() / sigma - 1.0) < 0.03 and abs(noise[inner].mean()) < 0.01
    out, ok = dec.decode_batch(llr, return_ok=True)
    exp, eok = ref.decode_batch(N, 1, fz, "char", llr.cpu().numpy())
    assert np.array_equal(out.cpu().numpy(), exp)
    counts = torch.zeros(4, dtype=torch.int64, device="cuda")
    pcb.capi.check(L.pc_count_errors(pcb.capi.ptr(info), pcb.capi.ptr(out), pcb.capi.ptr(ok), K, B, pcb.capi.ptr(counts), None))
    torch.cuda.synchronize()
    diff = np.unpackbits(h_info ^ exp, axis=1).sum(axis=1)
    assert counts.tolist() == [B, int((diff > 0).sum()), int(diff.sum()), int((eok == 0).sum())]
    assert 0.01 < counts[1].item() / B < 0.5


@pytest.mark.parametrize("N,K,B", [(1024, 512, 8192), (1024, 512, 40000), (1024, 512, 70000), (16384, 8192, 8192),
                                   (4096, 3072, 20000), (256, 128, 100000)])
def test_fastssc_small_batch_plans(pcb, ref, N, K, B):
    """batches that would leave SMs idle with the big-batch plan run a plan with smaller CTAs (pc_api_decode.inc
    run_fastssc_i8): same answers, checked against the reference on a slice and against the big-batch plan on all"""
    import torch
    rng = np.random.default_rng(B)
    fz = ref.frozen_bits(N, K, 0.0)
    dec = pcb.PolarDecoder(N, 1, list(fz), "char")
    base = awgn_llr_i8(rng, 2048, N, 2.0, K / N)
    llr = torch.from_numpy(base).cuda().repeat((B + 2047) // 2048, 1)[:B].contiguous()
    got, ok = dec.decode_batch(llr, return_ok=True)
    exp, eok = ref.decode_batch(N, 1, fz, "char", base, threads=8)
    g, o = got.cpu().numpy(), ok.cpu().numpy()
    for lo in range(0, B, 2048):
        n = min(2048, B - lo)
        assert np.array_equal(g[lo:lo + n], exp[:n]) and np.array_equal(o[lo:lo + n], eok[:n])


@pytest.mark.parametrize("N,K,B", [(1024, 512, 1 << 20), (16384, 8192, 1 << 16)])
def test_full_size_properties(pcb, N, K, B):
    """BASELINE.json batch sizes, size-independent properties of the code (no oracle needed): linearity of both encoder
    modes, the systematic property, and encode -> noiseless BPSK -> decode round trips through every decoder family"""
    import torch
    fz = pcb.frozen_bits(N, K, 0.0)
    kb = K // 8
    g = torch.Generator(device="cuda").manual_seed(N)
    a = torch.randint(0, 256, (B, kb), device="cuda", dtype=torch.uint8, generator=g)
    b = torch.randint(0, 256, (B, kb), device="cuda", dtype=torch.uint8, generator=g)
    enc = pcb.PolarEncoder(N, fz)                                   # Dummy detector: all K bits are payload
    info_pos = torch.ones(N, dtype=torch.bool, device="cuda")
    info_pos[torch.tensor(fz, device="cuda")] = False
    for syst in (True, False):
        enc.setSystematic(syst)
        ca, cb, cab = enc.encode_batch(a), enc.encode_batch(b), enc.encode_batch(a ^ b)
        assert torch.equal(ca ^ cb, cab)                            # linear code
        if syst:
            # systematic property on a slice (unpacking 2^20 x N bits at once would be 1 GiB of bools)
            n = min(B, 1 << 14)
            bits = ((ca[:n].unsqueeze(-1) >> torch.arange(7, -1, -1, device="cuda", dtype=torch.uint8)) & 1).reshape(n, N)
            ibits = ((a[:n].unsqueeze(-1) >> torch.arange(7, -1, -1, device="cuda", dtype=torch.uint8)) & 1).reshape(n, K)
            assert torch.equal(bits[:, info_pos], ibits)
    # round trips: transmitted word -> LLR +-100 (bit 0 -> +) -> decoder returns the info block, detector true
    enc.setSystematic(True)
    enc.setErrorDetection(8, "crc")
    code, sent = enc.encode_batch(a, return_info=True)              # `sent` carries the CRC-8 byte
    n = B if N == 1024 else B // 4
    for kind, L, frames in (("char", 1, n), ("float", 1, n // 4), ("char", 4, n // 64), ("float", 4, n // 128), (5, 4, n // 4)):
        dec = pcb.PolarDecoder(N, L, fz, kind)
        sub = code[:frames]
        bits = ((sub.unsqueeze(-1) >> torch.arange(7, -1, -1, device="cuda", dtype=torch.uint8)) & 1).reshape(frames, N)
        llr8 = (100 - 200 * bits.to(torch.int16)).to(torch.int8)
        llr = llr8 if kind in ("char", 5) else llr8.to(torch.float32)
        out, ok = dec.decode_batch(llr, return_ok=True)
        assert torch.equal(out, sent[:frames]) and bool((ok == 1).all())
        del bits, llr8, llr


def test_code_object_is_bound_to_its_device(pcb):
    """include/polar_b200.h "Devices, streams, alignment": a pc_code belongs to the device that was current at creation;
    with another device current every entry point answers PC_ERR_ARG instead of faulting (needs two GPUs)."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    N, K = 256, 128
    fz = pcb.frozen_bits(N, K, 0.0)
    with torch.cuda.device(0):
        dec = pcb.PolarDecoder(N, 1, fz, "char")
        enc = pcb.PolarEncoder(N, fz)
        llr0 = torch.zeros((8, N), dtype=torch.int8, device="cuda:0")
        assert dec.decode_batch(llr0).shape == (8, K // 8)
    with torch.cuda.device(1):
        llr1 = torch.zeros((8, N), dtype=torch.int8, device="cuda:1")
        with pytest.raises(ValueError, match="created on CUDA device 0"):
            dec.decode_batch(llr1)
        with pytest.raises(ValueError, match="created on CUDA device 0"):
            enc.encode_batch(torch.zeros((8, K // 8), dtype=torch.uint8, device="cuda:1"))
        # a second code object on the second device works, detector tables included (one table per device and kind)
        dec1 = pcb.PolarDecoder(N, 1, fz, "char")
        assert dec1.decode_batch(llr1).shape == (8, K // 8)
        det = pcb.Detector(16, "crc")
        assert det.check(det.generate(np.frombuffer(b"Test", np.uint8)))
    with torch.cuda.device(0):
        assert dec.decode_batch(llr0).shape == (8, K // 8)


def test_misaligned_pointers_are_an_argument_error(pcb):
    """16-byte aligned LLR batches, 4-byte aligned info batches: PC_ERR_ARG (ValueError), not a misaligned-address fault"""
    import torch
    N, K = 1024, 512
    fz = pcb.frozen_bits(N, K, 0.0)
    dec = pcb.PolarDecoder(N, 1, fz, "char")
    buf = torch.zeros(4 * N + 64, dtype=torch.int8, device="cuda")
    out = torch.empty((4, K // 8), dtype=torch.uint8, device="cuda")
    ok = torch.empty(4, dtype=torch.uint8, device="cuda")
    L_, ptr = pcb.capi.lib(), pcb.capi.ptr
    with pytest.raises(ValueError, match="16-byte aligned"):
        pcb.capi.check(L_.pc_decode_batch(dec._h, buf.data_ptr() + 4, 0, ptr(out), ptr(ok), 4, None))
    pcb.capi.check(L_.pc_decode_batch(dec._h, buf.data_ptr() + 16, 0, ptr(out), ptr(ok), 4, None))
    torch.cuda.synchronize()
