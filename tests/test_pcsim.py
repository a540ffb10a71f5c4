"""pcsim-compatible front-end (polar-codes_b200/pcsim.py): job lists, Eb/N0 inflation, decoder mapping and CSV schema
against the rules of the reference's simulator (src/simulation/setup.cpp, simulator.cpp; line numbers per test);
on the GPU a short run per simulation type whose CSV loads with the reference's own loader call and whose error
rates agree with the unmodified reference decoding the same device-generated LLRs."""
import os
import re

import numpy as np
import pytest


@pytest.fixture(scope="module")
def sim(pcb):
    return pcb.pcsim


def _cfg(sim, *argv):
    return sim.config_from_args(list(argv))


def test_defaults_match_setup_cpp(sim):
    """setup.cpp:25-66"""
    c = _cfg(sim)
    assert c["simtype"] == "single" and c["workload"] == 10 ** 9 and c["snr-count"] == 16
    assert c["precision"] == 832 and c["error-detection"] == "crc32" and c["amplification"] == 10.0
    assert c["blocklength"] == 1024 and c["rate"] == 0.5 and c["pathlimit"] == 8 and c["output"] == "simulation"
    dp = sim.default_data_point(c)
    assert (dp.N, dp.K, dp.L, dp.errorDetection, dp.errorDetectionType) == (1024, 512, 8, 32, "crc")
    assert dp.BlocksToSimulate == 10 ** 9 // 1024 and dp.systematic and dp.EbN0 == 4.0


def test_csv_header_is_the_reference_string(sim):
    """simulator.cpp:528-533 and :574-579: the literal is rebuilt here from the reference's pieces"""
    ref = ("\"N\",\"K\",\"dSNR\",\"C\",\"L\",\"Eb/" "N0\",\"BPS\",\"BLER\",\"BER\",\"RER\",\"Runs\",\"Errors\",\"Time\","
           "\"Blockspeed\",\"Coded Bitrate\",\"Payload Bitrate\",\"Effective Payload " "Bitrate\",\"Encoder Bitrate\","
           "\"Amplification\",\"time min\",\"time " "max\",\"time mean\",\"time deviation\"")
    assert sim.CSV_HEADER == ref
    assert sim.CSV_HEADER_COMPARE == '"Name",' + ref
    assert len(ref.split(",")) == 23


def test_snr_inflation_rule(sim):
    """simulator.cpp:359-402: ranges [snr-min,0], [0,2], [2,snr-max] with count/4, /2, /4 points, first point of each
    range skipped (loops start at i = 1), float arithmetic"""
    c = _cfg(sim)
    jobs = sim.build_job_list(c)
    assert len(jobs) == 3 + 7 + 3
    f = np.float32
    smin = f(-1.59174539)
    exp = [smin + f(i) * ((f(0) - smin) / f(3)) for i in range(1, 4)]
    exp += [f(0) + f(i) * (f(2) / f(7)) for i in range(1, 8)]
    exp += [f(2) + f(i) * (f(2) / f(3)) for i in range(1, 4)]
    assert [j.EbN0 for j in jobs] == [float(v) for v in exp]
    assert all(j.amplification == 10.0 for j in jobs)
    # precision 32: matched LLR scaling 4 * Eb/N0 (linear), simulator.cpp:366-379
    j32 = sim.build_job_list(_cfg(sim, "-p", "32"))
    for j in j32:
        assert j.amplification == float(f(4.0 * 10.0 ** (j.EbN0 / 10.0)))
    # tiny counts: the unsigned loops of the reference are empty
    assert sim.build_job_list(_cfg(sim, "--snr-count", "4")) == [] or len(sim.build_job_list(_cfg(sim, "--snr-count", "4"))) == 1


def test_job_lists_per_simtype(sim):
    """configure*Sim, simulator.cpp:164-269"""
    per = 13
    jobs = sim.build_job_list(_cfg(sim, "codelength", "--n-min", "128", "--n-max", "1024", "-r", "0.3", "-w", "1000000"))
    assert sorted({j.N for j in jobs}) == [128, 256, 512, 1024] and len(jobs) == 4 * per
    assert {(j.N, j.K) for j in jobs} == {(n, int(np.float32(n) * np.float32(0.3))) for n in (128, 256, 512, 1024)}
    assert all(j.BlocksToSimulate == 1000000 // j.N for j in jobs)
    jobs = sim.build_job_list(_cfg(sim, "listlength"))
    assert sorted({j.L for j in jobs}) == [1, 2, 4, 8, 16] and len(jobs) == 5 * per
    jobs = sim.build_job_list(_cfg(sim, "designsnr"))
    f = np.float32
    lo = f(-1.59174539)
    assert sorted({j.designSNR for j in jobs}) == sorted(float(lo + f(i) * ((f(6.0) - lo) / f(5))) for i in range(6))
    jobs = sim.build_job_list(_cfg(sim, "rate"))
    ks = []
    for i in range(5):
        k = int(f(1024) * (f(0.25) + f(i) * ((f(0.9) - f(0.25)) / f(4))))
        ks.append((k + 7) // 8 * 8)
    assert sorted({j.K for j in jobs}) == sorted(set(ks))
    jobs = sim.build_job_list(_cfg(sim, "amplification"))
    assert sorted({j.amplification for j in jobs}) == [float(f(1.0) + f(i) * (f(127.0) / f(5))) for i in range(6)]
    for st in ("fixed", "scan", "fastsscan", "ask"):
        with pytest.raises(NotImplementedError):
            sim.build_job_list(_cfg(sim, st))
    # depthfirst (simulator.cpp:303-321): the trial limit doubles from l-min to l-max, every job inflated over the SNR grid
    jobs = sim.build_job_list(_cfg(sim, "depthfirst", "--l-min", "2", "--l-max", "8"))
    assert sorted({j.L for j in jobs}) == [2, 4, 8] and {j.decoderType for j in jobs} == {"depthfirst"}
    assert sim.decoder_kind(jobs[0]) == sim.capi.PC_DEC_SCFLIP_F32


def test_compareall_jobs(sim):
    """simulator.cpp:404-491: (N, K, Eb/N0) = (1024, 512, 2 dB) and (4096, 3072, 3 dB) per decoder; no SNR inflation"""
    skipped = []
    jobs = sim.build_job_list(_cfg(sim, "compareall"), skipped)
    assert [j.name for j in jobs] == ["Fast-SSC32", "Fast-SSC32", "Fast-SSC8", "Fast-SSC8", "SCL32", "SCL32", "SCL8", "SCL8",
                                      "SCFlip", "SCFlip"]
    assert skipped == ["Fast-SSCAN", "Fast-SSCAN", "FFSSC"]
    assert [(j.N, j.K, j.EbN0) for j in jobs[:2]] == [(1024, 512, 2.0), (4096, 3072, 3.0)]
    assert [j.L for j in jobs] == [1, 1, 1, 1, 8, 8, 8, 8, 8, 8]
    assert [j.decoderType for j in jobs[8:]] == ["depthfirst"] * 2 and jobs[8].amplification == jobs[0].amplification
    assert jobs[0].amplification == float(np.float32(4 * 10 ** 0.2)) and jobs[1].amplification == float(np.float32(4 * 10 ** 0.3))
    assert jobs[2].amplification == 10.0 and jobs[7].amplification == 10.0


def test_decoder_mapping_and_detector_guard(sim, pcb):
    """setCoders simulator.cpp:703-765, setErrorDetector :766-819"""
    c = pcb.capi
    dp = sim.default_data_point(_cfg(sim))
    exp = {(8, 8): c.PC_DEC_ADAPTIVE_I8, (32, 8): c.PC_DEC_ADAPTIVE_F32, (832, 8): c.PC_DEC_ADAPTIVE_MIXED,
           (8, 1): c.PC_DEC_FASTSSC_I8, (832, 1): c.PC_DEC_FASTSSC_I8, (32, 1): c.PC_DEC_FASTSSC_F32}
    for (p, l), k in exp.items():
        dp.precision, dp.L = p, l
        assert sim.decoder_kind(dp) == k
    dp.precision = 16
    with pytest.raises(ValueError):
        sim.decoder_kind(dp)
    dp = sim.default_data_point(_cfg(sim, "-n", "32", "-r", "0.5", "-e", "crc32"))
    assert sim.effective_detector(dp) == (0, "none")              # errorDetection >= K
    dp = sim.default_data_point(_cfg(sim, "-e", "cmac16"))
    with pytest.raises(NotImplementedError):
        sim.effective_detector(dp)
    assert sim.effective_detector(sim.default_data_point(_cfg(sim, "-e", "crc8"))) == (8, "crc")
    assert sim.effective_detector(sim.default_data_point(_cfg(sim, "-e", "none"))) == (0, "none")


def test_csv_rows_load_with_the_reference_loader(sim, tmp_path):
    """saveResults formatting (simulator.cpp:535-557): %g floats, 1e-99 for zero rates, integer nanoseconds; the file
    is read back with the call python/analyze_results.py:164-166 makes"""
    c = _cfg(sim, "single", "-o", str(tmp_path / "res"))
    job = sim.default_data_point(c)
    job.runs, job.errors, job.biterrors, job.reportedErrors = 1000, 0, 0, 0
    sim.finish_job(job, 0.5, 0.25, np.array([1e-6, 3e-6]))
    job2 = job.copy()
    job2.errors, job2.biterrors, job2.reportedErrors, job2.EbN0 = 10, 123, 7, 1.5
    sim.finish_job(job2, 0.5, 0.25, np.array([1e-6, 3e-6]))
    name = sim.save_results([job, job2], c)
    assert name.endswith("res_single.csv")
    lines = open(name).read().splitlines()
    assert lines[0] == sim.CSV_HEADER
    cols = lines[1].split(",")
    assert cols[:7] == ["1024", "512", "0", "32", "8", "4", "1"] and cols[7:10] == ["1e-99"] * 3
    assert cols[10:12] == ["1000", "0"] and cols[12] == "0.5" and cols[13] == "2000" and cols[14] == "2.048e+06"
    assert cols[15] == "960000" and cols[16] == "960000" and cols[17] == "4.096e+06" and cols[18] == "10"
    assert cols[19:] == ["1000", "3000", "2000", "1000"]
    m = np.genfromtxt(name, delimiter=",", skip_header=1)
    assert m.shape == (2, 23)
    assert m[1, 7] == pytest.approx(0.01) and m[1, 8] == pytest.approx(123 / (1000 * 512), rel=1e-5) and m[1, 9] == pytest.approx(0.007)
    assert m[1, 16] == pytest.approx(990 * 480 / 0.5)


@pytest.mark.gpu
@pytest.mark.parametrize("argv", [("single", "-p", "8", "-e", "crc8", "-l", "4", "--snr-max", "2.5"),
                                  ("listlength", "-p", "832", "-e", "crc32", "--l-max", "4", "-n", "256"),
                                  ("codelength", "-p", "32", "-e", "crc8", "-l", "2", "--n-min", "64", "--n-max", "256"),
                                  ("rate", "-p", "8", "-e", "none", "-l", "1", "-n", "512", "--r-count", "3"),
                                  ("compareall", "-e", "crc8", "-l", "4")])
def test_pcsim_runs_and_writes_reference_csv(sim, tmp_path, argv):
    out = str(tmp_path / "gpu")
    rc = sim.main(list(argv) + ["-w", "600000", "--snr-count", "8", "-o", out, "--batch", "300"])
    assert rc == 0
    name = f"{out}_{argv[0]}.csv"
    lines = open(name).read().splitlines()
    compare = argv[0] == "compareall"
    assert lines[0] == (sim.CSV_HEADER_COMPARE if compare else sim.CSV_HEADER)
    m = np.genfromtxt(name, delimiter=",", skip_header=1)
    off = 1 if compare else 0
    assert m.ndim == 2 and m.shape[1] == 23 + off and m.shape[0] == len(lines) - 1 > 0
    N, K, runs, errors = m[:, off + 0], m[:, off + 1], m[:, off + 10], m[:, off + 11]
    # compareall copies the template's block count into its N=4096 rows (simulator.cpp:468-479)
    assert np.all(runs == 600000 // (1024 if compare else N))
    bler = m[:, off + 7]
    assert np.all((bler == 1e-99) | (np.abs(bler - errors / runs) < 1e-6))
    assert np.all(m[:, off + 12] > 0) and np.all(m[:, off + 15] > 0)       # decode time, payload bit rate
    if not compare:
        # error rate falls with Eb/N0 inside every parameter group (coarse: first vs last point)
        key = m[:, [0, 1, 2, 4]]
        for k in np.unique(key, axis=0):
            rows = m[(key == k).all(axis=1)]
            assert rows[0, 7] >= rows[-1, 7]


@pytest.mark.gpu
@pytest.mark.parametrize("prec,L", [(8, 1), (8, 4), (32, 4), (832, 4)])
def test_pcsim_counts_match_reference_on_the_same_llrs(sim, pcb, ref, prec, L):
    """the job's counters against the unmodified reference decoding the identical device-generated blocks"""
    import ctypes

    import torch
    c = _cfg(sim, "single", "-n", "256", "-p", str(prec), "-l", str(L), "-e", "crc8", "-w", str(256 * 2000))
    job = sim.default_data_point(c)
    job.EbN0 = 2.0
    if prec == 32:
        job.amplification = float(np.float32(4 * 10 ** 0.2))
    sim.run_job(job, seed=77, batch=2000, job_index=3)
    # regenerate the same blocks (same Philox counters) and decode them with the reference
    fz = pcb.frozen_bits(job.N, job.K, job.designSNR, "BB")
    dec = pcb.PolarDecoder(job.N, job.L, fz, sim.decoder_kind(job))
    dec.setErrorDetection(8, "crc")
    is_float = prec != 8
    llr = torch.empty((2000, job.N), dtype=torch.float32 if is_float else torch.int8, device="cuda")
    info = torch.empty((2000, job.K // 8), dtype=torch.uint8, device="cuda")
    st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    pcb.capi.check(pcb.capi.lib().pc_awgn_generate(dec._h, 77, (3 << 40), 2.0, job.amplification, int(is_float),
                                                   pcb.capi.ptr(llr), pcb.capi.ptr(info), 2000, st))
    torch.cuda.synchronize()
    h = llr.cpu().numpy()
    if not is_float:
        h = h.astype(np.float32)                                   # AdaptiveChar takes float LLRs only
    kind = {(8, 1): "char", (8, 4): "adaptive_char", (32, 4): "mixed", (832, 4): "adaptive_mixed"}[(prec, L)]
    exp, eok = ref.decode_batch(job.N, L, np.array(fz, np.uint32), kind, h, crc_bits=8, fresh=True, threads=8)
    sent = info.cpu().numpy()
    diff = np.unpackbits(exp ^ sent, axis=1).sum(axis=1)
    assert job.runs == 2000
    assert job.errors == int((diff > 0).sum())
    assert job.biterrors == int(diff.sum())
    assert job.reportedErrors == int((eok == 0).sum())
