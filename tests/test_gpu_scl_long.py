"""List decoders on long codes (N = 8192 ... 32768): the decision words of the largest stages move to the arena when the
path state of even one CTA does not fit shared memory (pcsim's codelength sweep runs to N = 32768 by default,
setup.cpp:39-41). Bit-exact against the unmodified reference."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _tx(ref, rng, N, K, fz, B, ebn0, amp):
    info = rng.integers(0, 256, (B, K // 8), dtype=np.uint8)
    code, _ = ref.encode_batch(N, fz, info, crc_bits=8, crc_type="crc")
    x = 1.0 - 2.0 * np.unpackbits(code, axis=1, bitorder="big")[:, :N].astype(np.float32)
    sigma = 1.0 / np.sqrt(2.0 * (K / N) * 10.0 ** (ebn0 / 10.0))
    return ((x + sigma * rng.standard_normal((B, N))) * amp).astype(np.float32)


@pytest.mark.parametrize("N,K,L,kind", [(8192, 4096, 8, "char"), (16384, 8192, 8, "char"), (32768, 16384, 4, "char"),
                                        (16384, 8192, 32, "char"), (16384, 8192, 8, "float"), (32768, 16384, 8, "float"),
                                        (16384, 8192, 8, 5), (32768, 16384, 8, 6)])
def test_list_decoders_long_codes(pcb, ref, N, K, L, kind):
    rng = np.random.default_rng(N + L)
    fz = ref.frozen_bits(N, K, 0.0)
    dec = pcb.PolarDecoder(N, L, list(fz), kind)
    dec.setErrorDetection(8, "crc")
    B = 24 if N <= 16384 else 12
    # low Eb/N0 so that the first pass of the adaptive kinds fails on part of the batch
    y = np.concatenate([_tx(ref, rng, N, K, fz, B // 2, 0.9, 1.0), _tx(ref, rng, N, K, fz, B - B // 2, 1.6, 1.0)])
    if kind == "char":
        llr = np.clip(np.rint(10.0 * y), -128, 127).astype(np.int8)
        refkind, fresh = "char", False
    elif kind == "float":
        llr, refkind, fresh = (4.0 * y).astype(np.float32), "float", True
    else:
        llr, refkind, fresh = (10.0 * y).astype(np.float32), {5: "adaptive_char", 6: "adaptive_mixed"}[kind], kind == 6
    got, ok = dec.decode_batch(llr, return_ok=True)
    exp, eok = ref.decode_batch(N, L, fz, refkind, llr, crc_bits=8, fresh=fresh, threads=12)
    assert np.array_equal(got, exp) and np.array_equal(ok, eok)
    assert 0 < ok.mean()
