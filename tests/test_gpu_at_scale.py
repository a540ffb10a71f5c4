"""At-scale parity gate of SURVEY.md 8(d) inside `pytest -m gpu`: frames from the device-side channel generator at
BASELINE.json's batch sizes are decoded on the GPU, the SAME LLRs are copied back and replayed through the unmodified
reference (oracle/_ref, all host threads), info bytes and detector verdicts are compared frame by frame.

Bars: int8 decoders, encoder: 0 mismatching frames / verdicts. fp32 list decoder: identical on >= 99.99 % of the frames
(north_star's tolerance; asserted at 100 % where it holds and reported). Host cost: < 10 s in total on 16 threads.
"""
import ctypes
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _generate(pcb, dec, torch, frames, ebn0, is_float, amp, kb, N, first=0):
    L_, ptr, check = pcb.capi.lib(), pcb.capi.ptr, pcb.capi.check
    llr = torch.empty((frames, N), dtype=torch.float32 if is_float else torch.int8, device="cuda")
    info = torch.empty((frames, kb), dtype=torch.uint8, device="cuda")
    check(L_.pc_awgn_generate(dec._h, ctypes.c_uint64(0xA75CA1E), ctypes.c_uint64(first), ctypes.c_float(ebn0),
                              ctypes.c_float(amp), int(is_float), ptr(llr), ptr(info), frames, None))
    return llr, info


def _compare(pcb, ref, N, K, L, typ, crc, frames, ebn0, is_float=False, fresh=False, first=0):
    import torch
    fz = pcb.frozen_bits(N, K, 0.0)
    dec = pcb.PolarDecoder(N, L, fz, typ)
    dec.setErrorDetection(crc, "crc")
    kb = (K + 7) // 8
    amp = 4.0 * 10.0 ** (ebn0 / 10.0) if is_float else 10.0
    llr, info = _generate(pcb, dec, torch, frames, ebn0, is_float, amp, kb, N, first)
    out, ok = dec.decode_batch(llr, return_ok=True)
    h_llr, h_out, h_ok = llr.cpu().numpy(), out.cpu().numpy(), ok.cpu().numpy()
    del llr, out, ok
    exp, eok = ref.decode_batch(N, L, np.array(fz, np.uint32), typ, h_llr, crc_bits=crc, crc_type="crc",
                                threads=os.cpu_count() or 1, fresh=fresh)
    bad = int((h_out != exp).any(axis=1).sum())
    badv = int((h_ok != eok).sum())
    bler = float((exp != info.cpu().numpy()).any(axis=1).mean())
    return bad, badv, bler


@pytest.mark.parametrize("ebn0", [1.0, 2.0, 3.0])
def test_fastssc_i8_2e20_frames(pcb, ref, ebn0):
    """C2 at BASELINE's batch: 2^20 frames per point, FastSscFipChar replayed on the same LLRs"""
    bad, badv, bler = _compare(pcb, ref, 1024, 512, 1, "char", 8, 1 << 20, ebn0, first=int(ebn0 * 10) << 32)
    assert (bad, badv) == (0, 0), f"{bad} frames / {badv} verdicts differ from the reference at {ebn0} dB"
    assert 0.0 < bler < 1.0 or ebn0 >= 3.0


def test_scl_i8_l8_2e15_frames(pcb, ref):
    """the metric's second half: CRC-16 SCL L=8 int8 against SclFipChar, survivor order and ties included"""
    bad, badv, _ = _compare(pcb, ref, 1024, 512, 8, "char", 16, 1 << 15, 2.0)
    assert (bad, badv) == (0, 0)


def test_scl_i8_l32_2e13_frames(pcb, ref):
    """C4: list size 32 (stress of list copy and top-L selection)"""
    bad, badv, _ = _compare(pcb, ref, 1024, 512, 32, "char", 16, 1 << 13, 1.5)
    assert (bad, badv) == (0, 0)


def test_scl_f32_l8_2e14_frames(pcb, ref):
    """SclAvxFloat (fresh reference object per frame). Tolerance: >= 99.99 % identical frames."""
    frames = 1 << 14
    bad, badv, _ = _compare(pcb, ref, 1024, 512, 8, "float", 16, frames, 2.0, is_float=True, fresh=True)
    assert bad <= frames * 1e-4 and badv <= frames * 1e-4, f"{bad} / {badv} of {frames} differ"
    assert bad == 0 and badv == 0  # holds bit for bit today; the bar above is the contract


def test_c3_scl_f32_crc24_2e14_frames(pcb, ref):
    """C3's decoder: N=512 K=128 CRC-24C CA-SCL L=8 float"""
    frames = 1 << 14
    bad, badv, _ = _compare(pcb, ref, 512, 128, 8, "float", 24, frames, 1.0, is_float=True, fresh=True)
    assert bad <= frames * 1e-4 and badv <= frames * 1e-4


def test_fastssc_f32_2e18_frames(pcb, ref):
    bad, badv, _ = _compare(pcb, ref, 1024, 512, 1, "float", 8, 1 << 18, 2.0, is_float=True)
    assert bad <= (1 << 18) * 1e-4 and badv <= (1 << 18) * 1e-4
    assert bad == 0 and badv == 0


def test_c5_fastssc_i8_2e16_frames(pcb, ref):
    """C5: N=16384 K=8192 Fast-SSC int8 at BASELINE's 2^16 frames per launch"""
    bad, badv, _ = _compare(pcb, ref, 16384, 8192, 1, "char", 8, 1 << 16, 2.0)
    assert (bad, badv) == (0, 0)


@pytest.mark.parametrize("N,K,rows,crc", [(16384, 8192, 1 << 16, 8), (1024, 512, 1 << 18, 16), (16384, 8192, 1 << 14, 0)])
@pytest.mark.parametrize("systematic", [True, False])
def test_encoder_at_scale(pcb, ref, N, K, rows, crc, systematic):
    """C5's encoder batch diffed against ButterflyFipPacked row by row (check bytes included)"""
    if not systematic and rows > (1 << 14):
        rows = 1 << 14
    rng = np.random.default_rng(N + crc)
    fz = pcb.frozen_bits(N, K, 0.0)
    info = rng.integers(0, 256, (rows, K // 8), dtype=np.uint8)
    enc = pcb.PolarEncoder(N, fz)
    enc.setErrorDetection(crc, "crc")
    enc.setSystematic(systematic)
    got = enc.encode_batch(info)
    got = got.cpu().numpy() if hasattr(got, "cpu") else np.asarray(got)
    exp, _ = ref.encode_batch(N, np.array(fz, np.uint32), info, crc_bits=crc, systematic=systematic,
                              threads=os.cpu_count() or 1)
    assert np.array_equal(got, exp), f"{int((got != exp).any(axis=1).sum())} code words differ"


@pytest.mark.parametrize("crc,typ", [(6, "crcnr"), (8, "crc"), (11, "crcnr"), (16, "crc"), (16, "crcnr"), (24, "crc"), (32, "crc")])
@pytest.mark.parametrize("N,K", [(1024, 512), (256, 128), (64, 40), (512, 264), (32, 16)])
def test_encoder_fused_detector(pcb, ref, crc, typ, N, K):
    """pc_encode_batch without the info-copy output: for N <= 1024 the detector's generate runs INSIDE the encoder kernel
    on the staged rows (pc_encode2.cu). Same code words as the reference's generate-then-encode."""
    import torch
    kb = (K + 7) // 8
    cb = {6: 1, 8: 1, 11: 2, 16: 2, 24: 3, 32: 4}[crc]
    if crc == 32 and kb % 4:
        pytest.skip("CRC-32C needs whole words (crc32.cpp)")
    if kb <= cb:
        pytest.skip("block shorter than the checksum")
    rng = np.random.default_rng(N + crc)
    fz = pcb.frozen_bits(N, K, 0.0)
    B = 3000
    info = rng.integers(0, 256, (B, kb), dtype=np.uint8)
    for systematic in (True, False):
        enc = pcb.PolarEncoder(N, fz)
        enc.setErrorDetection(crc, typ)
        enc.setSystematic(systematic)
        d_info = torch.from_numpy(info).cuda()
        d_code = torch.empty((B, N // 8), dtype=torch.uint8, device="cuda")
        L_, ptr, check = pcb.capi.lib(), pcb.capi.ptr, pcb.capi.check
        check(L_.pc_encode_batch(enc._h, ptr(d_info), ptr(d_code), None, B, None))
        torch.cuda.synchronize()
        exp, _ = ref.encode_batch(N, np.array(fz, np.uint32), info, crc_bits=crc, crc_type=typ, systematic=systematic,
                                  threads=os.cpu_count() or 1)
        assert np.array_equal(d_code.cpu().numpy(), exp), f"systematic={systematic}"
        assert np.array_equal(d_info.cpu().numpy(), info)  # the caller's blocks stay untouched
