import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def pcb():
    import polar_b200_loader
    return polar_b200_loader.load()


@pytest.fixture(scope="session")
def ref():
    """the unmodified reference behind oracle/_ref/libref_shim.so (test oracle only)"""
    import refshim
    if not refshim.available():
        pytest.skip("oracle/_ref not built (run `make -C oracle`)")
    return refshim


@pytest.fixture(scope="session")
def orc():
    """our scalar C restatement oracle/libpc_oracle.so (test oracle only)"""
    import pcoracle
    return pcoracle


def awgn_llr_i8(rng, B, N, ebn0_db, rate, amp=10.0, codewords=None):
    """BPSK (0 -> +1) over AWGN, LLR = sat(rint(amp * y)) like pcsim"""
    sigma = 1.0 / np.sqrt(2.0 * rate * 10.0 ** (ebn0_db / 10.0))
    x = np.ones((B, N), dtype=np.float32) if codewords is None else (1.0 - 2.0 * codewords.astype(np.float32))
    y = x + sigma * rng.standard_normal((B, N), dtype=np.float32)
    return np.clip(np.rint(amp * y), -128, 127).astype(np.int8)


def unpack_bits(bytes_arr, nbits):
    return np.unpackbits(bytes_arr, axis=-1, bitorder="big")[..., :nbits]
