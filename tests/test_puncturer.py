"""Puncturer: batch gather / scatter kernels behind pc_puncture*_batch, the pypolar-compatible
`Puncturer` class and the C++ mirror, against the unmodified reference (oracle/ref_shim.cpp ref_puncturer) and in the
shape of the reference's own python/qa_pypolar_puncturer.py."""
import numpy as np
import pytest


def _positions(N, punctured):
    return np.setdiff1d(np.arange(N), punctured)


@pytest.mark.gpu
def test_setup_like_reference_qa(pcb, ref):
    """qa_pypolar_puncturer.py:30-53"""
    for n in range(5, 11):
        N = 2 ** n
        f = ref.frozen_bits(N, N // 2, 0.0)
        for cut in (N // 4, N // 8):
            p = pcb.Puncturer(N - cut, f)
            assert p.parentBlockLength() == N and p.blockLength() == N - cut
            assert p.blockOutputPositions() == _positions(N, f[:cut]).tolist()
            assert p.blockOutputPositions() == ref.puncturer(N - cut, f, "positions").tolist()


@pytest.mark.gpu
def test_puncture_and_depuncture_vectors_like_reference_qa(pcb, ref):
    """qa_pypolar_puncturer.py:55-112"""
    N = 64
    rng = np.random.default_rng(3)
    f = ref.frozen_bits(N, N // 2, 0.0)
    pos = _positions(N, f[:N // 4])
    p = pcb.Puncturer(N - N // 4, f)
    vec = rng.integers(0, 256, N // 8, dtype=np.uint8)
    unp = np.unpackbits(vec)
    res_bits = p.puncture(unp)
    res_packed = p.puncturePacked(vec)
    assert res_bits.dtype == np.uint8 and np.unpackbits(res_packed).tolist() == res_bits.tolist() == unp[pos].tolist()
    for dt in (np.float32, np.float64):
        v = np.arange(N, dtype=dt)
        r = p.puncture(v)
        assert r.dtype == dt and r.tolist() == v[pos].tolist()
        w = rng.normal(0.0, 1.0, N - N // 4).astype(dt)
        d = p.depuncture(w)
        e = np.zeros(N, dtype=dt)
        e[pos] = w
        assert d.dtype == dt and d.tolist() == e.tolist()
    w = rng.integers(0, 256, N - N // 4, dtype=np.uint8)
    e = np.zeros(N, dtype=np.uint8)
    e[pos] = w
    assert p.depuncture(w).tolist() == e.tolist()
    # argument checks of the bindings (puncturer_python.cc:33-38, 50-55, 102-107)
    with pytest.raises(RuntimeError, match="ONE-dimensional"):
        p.puncture(np.zeros((2, N), np.float32))
    with pytest.raises(RuntimeError, match="parentBlockSize"):
        p.puncture(np.zeros(N - 1, np.float32))
    with pytest.raises(RuntimeError, match="blockSize"):
        p.depuncture(np.zeros(N, np.float32))
    with pytest.raises(IndexError):                              # std::out_of_range, puncturer.cpp:57-60
        pcb.Puncturer(40, f[:10])


@pytest.mark.gpu
@pytest.mark.parametrize("M,N,K", [(48, 64, 32), (56, 64, 16), (768, 1024, 512), (1000, 1024, 256), (520, 1024, 512),
                                   (12288, 16384, 8192), (1024, 1024, 512), (24, 32, 16)])
def test_batch_kernels_match_reference(pcb, ref, M, N, K):
    import torch
    rng = np.random.default_rng(M)
    f = ref.frozen_bits(N, K, 0.0)
    p = pcb.Puncturer(M, f)
    B = 257 if N <= 1024 else 9
    for dt in (np.uint8, np.float32, np.float64):
        x = (rng.standard_normal((B, N)) * 50).astype(dt)
        assert np.array_equal(p.puncture_batch(x), ref.puncturer(M, f, "puncture", x))
        y = (rng.standard_normal((B, M)) * 50).astype(dt)
        assert np.array_equal(p.depuncture_batch(y), ref.puncturer(M, f, "depuncture", y))
    if M % 8 == 0:
        bits = rng.integers(0, 256, (B, N // 8), dtype=np.uint8)
        assert np.array_equal(p.puncturePacked_batch(bits), ref.puncturer(M, f, "packed", bits))
    # int8 LLRs on the device: depuncture of a puncture keeps the surviving positions and zeroes (erases) the rest
    llr = torch.from_numpy(rng.integers(-128, 128, (B, N), dtype=np.int8)).cuda()
    rt = p.depuncture_batch(p.puncture_batch(llr))
    keep = torch.zeros(N, dtype=torch.bool, device="cuda")
    keep[torch.tensor(p.blockOutputPositions(), device="cuda")] = True
    assert torch.equal(rt[:, keep], llr[:, keep]) and int(rt[:, ~keep].abs().sum()) == 0
    assert p.puncture_batch(np.zeros((0, N), np.float32)).shape == (0, M)


@pytest.mark.gpu
def test_punctured_code_round_trip(pcb, ref):
    """rate matching chain: encode (N=1024) -> puncture to 800 -> BPSK/AWGN -> depuncture (LLR 0 = erased) -> decode"""
    N, K, M = 1024, 400, 800
    rng = np.random.default_rng(9)
    f = ref.frozen_bits(N, K, 0.0)
    p = pcb.Puncturer(M, f)
    enc = pcb.PolarEncoder(N, f)
    enc.setErrorDetection(8, "crc")
    dec = pcb.PolarDecoder(N, 1, f, "float")
    info = rng.integers(0, 256, (300, K // 8), dtype=np.uint8)
    code = enc.encode_batch(info)
    tx = p.puncturePacked_batch(code)
    y = (1.0 - 2.0 * np.unpackbits(tx, axis=1).astype(np.float32)) + 0.45 * rng.standard_normal((300, M)).astype(np.float32)
    llr = p.depuncture_batch((4.0 * y).astype(np.float32))
    got, ok = dec.decode_batch(llr, return_ok=True)
    exp, eok = ref.decode_batch(N, 1, f, "float", llr, crc_bits=8)
    assert np.array_equal(got, exp) and np.array_equal(ok, eok)
    assert ok.mean() > 0.9 and np.array_equal(got[ok == 1][:, :K // 8 - 1], info[ok == 1][:, :K // 8 - 1])


def test_puncturer_without_gpu_fails_loudly(pcb):
    """no CPU path: the constructor refuses when no device is usable"""
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(RuntimeError, match="no usable CUDA device"):
        pcb.Puncturer(48, list(range(32)))
