"""SC-Flip: PC_DEC_SCFLIP_F32 against the reference's DepthFirst decoder
(src/polarcode/decoding/depth_first.cpp:556-640), the decoder pcsim builds for its `depthfirst` simulation type with the
list size as trial limit (simulator.cpp:715-716). Bar: identical info bytes and detector verdicts on the same float LLRs
(the leaf LLRs, their ranking and the configuration queue are reproduced exactly; see pc_scflip_f32.cu for the tie rule)."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _frames(ref, rng, N, K, B, ebn0, crc):
    fz = ref.frozen_bits(N, K, 0.0)
    info = rng.integers(0, 256, (B, K // 8), dtype=np.uint8)
    code, sent = ref.encode_batch(N, fz, info, crc_bits=crc, threads=os.cpu_count() or 1)
    sigma = 1.0 / np.sqrt(2.0 * (K / N) * 10.0 ** (ebn0 / 10.0))
    y = (1.0 - 2.0 * np.unpackbits(code, axis=1).astype(np.float32)) + np.float32(sigma) * rng.standard_normal((B, N)).astype(np.float32)
    return fz, sent, (np.float32(4.0 * 10.0 ** (ebn0 / 10.0)) * y).astype(np.float32)


@pytest.mark.parametrize("N,K,T,ebn0,crc", [(64, 32, 4, 1.0, 8), (128, 64, 8, 1.5, 8), (256, 128, 8, 1.5, 8), (1024, 512, 8, 2.0, 8),
                                            (1024, 512, 32, 1.5, 32), (512, 128, 16, 0.5, 8), (64, 40, 2, 0.5, 8),
                                            (1024, 768, 8, 3.0, 8), (2048, 1024, 4, 2.0, 8), (1024, 512, 1, 2.0, 8)])
def test_scflip_matches_reference(pcb, ref, N, K, T, ebn0, crc):
    rng = np.random.default_rng(N + K + T)
    B = 4096 if N <= 1024 else 1024
    fz, sent, llr = _frames(ref, rng, N, K, B, ebn0, crc)
    dec = pcb.PolarDecoder(N, T, list(fz), pcb.capi.PC_DEC_SCFLIP_F32)
    dec.setErrorDetection(crc, "crc")
    got, ok = dec.decode_batch(llr, return_ok=True)
    exp, eok = ref.decode_batch(N, T, fz, "depthfirst", llr, crc_bits=crc, threads=os.cpu_count() or 1)
    bad = int(((got != exp).any(axis=1) | (ok != eok)).sum())
    assert bad == 0, f"{bad} of {B} frames differ from DepthFirst"
    if T > 1:  # the flips must actually rescue frames, otherwise the test would not exercise the configuration queue
        one = pcb.PolarDecoder(N, 1, list(fz), pcb.capi.PC_DEC_SCFLIP_F32)
        one.setErrorDetection(crc, "crc")
        _, ok1 = one.decode_batch(llr, return_ok=True)
        assert ok.sum() > ok1.sum()
    good = ok == 1
    assert (got[good] == sent[good]).all(axis=1).mean() > 0.95  # (CRC-8 lets a few wrong words through)


def test_scflip_2e16_frames_three_snr_points(pcb, ref):
    """identical info bytes and verdicts on 2^16 frames at 1 - 3 dB (N=1024, K=512, trial limit 8)"""
    N, K, T = 1024, 512, 8
    for ebn0 in (1.0, 2.0, 3.0):
        rng = np.random.default_rng(int(ebn0 * 10))
        B = 1 << 16
        fz, _, llr = _frames(ref, rng, N, K, B, ebn0, 8)
        dec = pcb.PolarDecoder(N, T, list(fz), pcb.capi.PC_DEC_SCFLIP_F32)
        got, ok = dec.decode_batch(llr, return_ok=True)
        exp, eok = ref.decode_batch(N, T, fz, "depthfirst", llr, crc_bits=8, threads=os.cpu_count() or 1)
        assert int(((got != exp).any(axis=1) | (ok != eok)).sum()) == 0, f"mismatch at {ebn0} dB"


def test_scflip_non_systematic_and_int8_input(pcb, ref):
    N, K, T = 256, 128, 8
    rng = np.random.default_rng(11)
    fz, _, llr = _frames(ref, rng, N, K, 2000, 1.5, 8)
    dec = pcb.PolarDecoder(N, T, list(fz), pcb.capi.PC_DEC_SCFLIP_F32)
    dec.setSystematic(False)
    got, ok = dec.decode_batch(llr, return_ok=True)
    exp, eok = ref.decode_batch(N, T, fz, "depthfirst", llr, crc_bits=8, systematic=False)
    assert np.array_equal(got, exp) and np.array_equal(ok, eok)
