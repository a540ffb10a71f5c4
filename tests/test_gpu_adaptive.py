"""GPU parity of the two wrappers pcsim constructs by class (reference src/simulation/simulator.cpp:721-733):
AdaptiveChar (FastSscFipChar, then SclFipChar on detector failure; adaptive_char.cpp:35-46) and AdaptiveMixed
(FastSscFipChar, then SclAvxFloat; adaptive_mixed.cpp:36-47). Both take float LLRs in the reference
(setSignal(const float*) feeds the quantiser of the char decoder and the float container of the list decoder)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

CASES = [(1024, 512, 8, 16, "crc"), (256, 128, 4, 8, "crc"), (64, 40, 2, 8, "crc"), (512, 128, 8, 16, "crcnr"),
         (2048, 1024, 4, 24, "crc"), (128, 64, 32, 8, "crc"), (1024, 768, 8, 32, "crc"), (16, 8, 4, 8, "crc")]


def _tx(ref, rng, N, K, fz, cb, ct, B, amp):
    info = rng.integers(0, 256, (B, (K + 7) // 8), dtype=np.uint8)
    code, _ = ref.encode_batch(N, fz, info, crc_bits=cb, crc_type=ct)
    x = 1.0 - 2.0 * np.unpackbits(code, axis=1, bitorder="big")[:, :N].astype(np.float32)
    out = []
    for ebn0 in (0.5, 2.0, 3.5):
        sigma = 1.0 / np.sqrt(2.0 * (K / N) * 10.0 ** (ebn0 / 10.0))
        out.append(((x + sigma * rng.standard_normal((B, N))) * amp).astype(np.float32))
    return np.concatenate(out)


@pytest.mark.parametrize("N,K,L,cb,ct", CASES)
@pytest.mark.parametrize("kind,refkind", [(5, "adaptive_char"), (6, "adaptive_mixed")])
def test_pcsim_adaptive_parity(pcb, ref, kind, refkind, N, K, L, cb, ct):
    rng = np.random.default_rng(N + L + kind)
    fz = ref.frozen_bits(N, K, 0.0)
    dec = pcb.PolarDecoder(N, L, list(fz), kind)
    dec.setErrorDetection(cb, ct)
    llr = _tx(ref, rng, N, K, fz, cb, ct, 150 if N <= 1024 else 40, 10.0)
    got, ok = dec.decode_batch(llr, return_ok=True)
    exp, eok = ref.decode_batch(N, L, fz, refkind, llr, crc_bits=cb, crc_type=ct, fresh=True, threads=8)
    assert (got != exp).any(axis=1).mean() <= 1e-4          # north_star bar for the float list decoder
    assert np.array_equal(got, exp) and np.array_equal(ok, eok)
    if N in (64, 256, 1024):
        assert 0 < ok.mean() <= 1
    # the second pass really ran: a plain Fast-SSC pass leaves more failures
    fast = pcb.PolarDecoder(N, 1, list(fz), "char")
    fast.setErrorDetection(cb, ct)
    _, fok = fast.decode_batch(llr, return_ok=True)
    assert ok.sum() >= fok.sum()
    assert np.array_equal(got[fok == 1], fast.decode_batch(llr)[fok == 1])


def test_adaptive_char_int8_input_and_device_tensors(pcb, ref):
    """int8 batches are taken as already quantised (the reference wrapper has no char entry: decoder.cpp:140 would
    dereference a null container); device tensors run the same three launches on the caller's stream"""
    import torch
    N, K, L = 1024, 512, 8
    rng = np.random.default_rng(5)
    fz = ref.frozen_bits(N, K, 0.0)
    llr = _tx(ref, rng, N, K, fz, 16, "crc", 300, 10.0)
    q = ref.quantize(N, llr)
    dec = pcb.PolarDecoder(N, L, list(fz), 5)
    dec.setErrorDetection(16, "crc")
    a, oka = dec.decode_batch(llr, return_ok=True)
    b, okb = dec.decode_batch(q, return_ok=True)
    assert np.array_equal(a, b) and np.array_equal(oka, okb)
    t, okt = dec.decode_batch(torch.from_numpy(q).cuda(), return_ok=True)
    assert np.array_equal(t.cpu().numpy(), a) and np.array_equal(okt.cpu().numpy(), oka)
    # L = 1 degenerates to the Fast-SSC pass alone (adaptive_char.cpp:40 `mListSize > 1`)
    d1 = pcb.PolarDecoder(N, 1, list(fz), 5)
    d1.setErrorDetection(16, "crc")
    f = pcb.PolarDecoder(N, 1, list(fz), "char")
    f.setErrorDetection(16, "crc")
    assert np.array_equal(d1.decode_batch(q), f.decode_batch(q))
