"""GPU parity against the committed reference outputs for the list and float decoders
(tests/golden/reference_vectors_list.npz, written by tests/golden/make_golden_list.py from the unmodified reference).
Needs neither /root/reference nor oracle/_ref at run time."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

GOLD = np.load(os.path.join(os.path.dirname(__file__), "golden", "reference_vectors_list.npz"))
KEYS = sorted(k[:-4] for k in GOLD.files if k.endswith("_llr"))
KIND = {"sclc": "char", "float": "float", "mixed": "mixed", "adaptivechar": 5, "adaptivemixed": 6}   # 5/6: PC_DEC_ADAPTIVE_I8/_MIXED


@pytest.mark.parametrize("key", KEYS)
def test_list_and_float_decoders_golden(pcb, key):
    tag, N, K, L, cb, ct = key.split("_")
    N, L, cb = int(N), int(L), int(cb)
    dec = pcb.PolarDecoder(N, L, list(GOLD[key + "_frozen"]), KIND[tag])
    dec.setErrorDetection(cb, ct)
    llr = GOLD[key + "_llr"]
    got, ok = dec.decode_batch(llr, return_ok=True)
    assert np.array_equal(got, GOLD[key + "_info"])          # bit-exact; north_star asks >= 99.99 % for float
    assert np.array_equal(ok, GOLD[key + "_ok"])
    # ragged: one frame at a time gives the same answer as the batch (no cross-frame state on the device)
    for i in (0, llr.shape[0] - 1):
        g1, o1 = dec.decode_batch(llr[i:i + 1], return_ok=True)
        assert np.array_equal(g1[0], got[i]) and o1[0] == ok[i]
