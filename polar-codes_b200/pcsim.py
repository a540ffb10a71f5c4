"""pcsim-compatible sweep front-end: the reference's BLER simulator
(src/simulation/{setup,simulator}.cpp) with every per-block step on the GPU.

Same simulation types, options, defaults, job lists, Eb/N0 inflation rule and CSV schema as the reference, so that
python/analyze_results.py (`np.genfromtxt(file, delimiter=',', skip_header=1)`, analyze_results.py:164-166) and
plot_polar_code_benchmark.py keep working on the files written here:

    python tools/pcsim.py listlength -n 1024 -r 0.5 -e crc8 -p 8 -w 100000000 -o gpu

  reference                                           here
  --------------------------------------------------  -------------------------------------------------------------
  Configurator defaults        setup.cpp:25-66        DEFAULTS
  Simulator::Simulator         simulator.cpp:41-78    build_job_list()
  configure*Sim                simulator.cpp:164-269  _configure_*()
  pushJobsInRange / inflate    simulator.cpp:359-402  snr_inflate()  (float arithmetic, same point set)
  configureComparisonSim       simulator.cpp:404-491  _configure_compareall() (jobs whose decoder exists on the GPU)
  SimulationWorker::setCoders  simulator.cpp:703-765  decoder_kind()
  SimulationWorker::run        simulator.cpp:632-672  run_job(): pc_awgn_generate -> pc_decode_batch -> pc_count_errors
  calculateStatistics          simulator.cpp:967-985  finish_job()
  saveResults                  simulator.cpp:519-563  save_results()

Differences, all forced by batching: timing columns are CUDA-event times of the batched launches (per-block time =
batch time / frames in the batch; min / max / deviation are over batches), "Encoder Bitrate" is measured on the fused
generator launch (random info, detector, encode, BPSK, AWGN, scaling), `--threads` is accepted and ignored, frames are
sharded over the ranks of an initialised torch.distributed group with one all_reduce(sum) of the counters per job.
Out of scope (no GPU kernel, no CPU fallback): cmac detectors, the ask / fixed / scan / fastsscan simulation types, and
the Fast-SSCAN / FFSSC rows of compareall (skipped with a note on stderr). `depthfirst` (SC-Flip, DepthFirst with the list
size as trial limit, simulator.cpp:303-321,715-716) runs on PC_DEC_SCFLIP_F32."""
import argparse
import ctypes
import dataclasses
import math
import sys

import numpy as np

from . import capi
from .capi import check, lib, ptr

f32 = np.float32

# setup.cpp:25-66
DEFAULTS = {
    "simtype": "single", "workload": int(1e9), "snr-min": -1.59174539, "snr-max": 4.0, "snr-count": 16,
    "design-snr": 0.0, "dsnr-min": -1.59174539, "dsnr-max": 6.0, "dsnr-count": 6,
    "blocklength": 1024, "n-min": 128, "n-max": 32768,
    "rate": 0.5, "r-min": 0.25, "r-max": 0.9, "r-count": 5,
    "pathlimit": 8, "l-min": 1, "l-max": 16,
    "error-detection": "crc32", "non-systematic": False, "precision": 832,
    "amplification": 10.0, "amp-min": 1.0, "amp-max": 128.0, "amp-count": 6,
    "output": "simulation", "threads": 1,
}
SIM_TYPES = ["single", "codelength", "designsnr", "listlength", "rate", "amplification", "fixed", "depthfirst", "scan",
             "fastsscan", "ask", "compareall", "getcode"]                                   # setup.cpp:77-81
GPU_SIM_TYPES = ["single", "codelength", "designsnr", "listlength", "rate", "amplification", "depthfirst", "compareall",
                 "getcode"]
ERROR_DETECTIONS = ["none", "crc8", "crc32", "cmac8", "cmac16", "cmac32", "cmac64", "cmac128"]  # setup.cpp:226-228

# saveResults, simulator.cpp:528-533
CSV_HEADER = ('"N","K","dSNR","C","L","Eb/N0","BPS","BLER","BER","RER","Runs","Errors","Time","Blockspeed",'
              '"Coded Bitrate","Payload Bitrate","Effective Payload Bitrate","Encoder Bitrate","Amplification",'
              '"time min","time max","time mean","time deviation"')
CSV_HEADER_COMPARE = '"Name",' + CSV_HEADER                                                # simulator.cpp:574-579


def error_detection_id(s):
    """errorDetectionStringToId, simulator.h:33-53"""
    return {"crc8": 8, "crc32": 32, "cmac8": 8, "cmac16": 16, "cmac32": 32, "cmac64": 64, "cmac128": 128}.get(s, 0)


def error_detection_type(s):
    """errorDetectionStringToType, simulator.h:56-67"""
    return "crc" if s.startswith("crc") else ("cmac" if s.startswith("cmac") else "none")


@dataclasses.dataclass
class DataPoint:
    """struct DataPoint, simulator.h:75-122 (the fields the GPU path uses)"""
    name: str = ""
    designSNR: float = 0.0
    N: int = 1024
    K: int = 512
    L: int = 8
    errorDetection: int = 32
    errorDetectionType: str = "crc"
    systematic: bool = True
    decoderType: str = "flexible"
    EbN0: float = 4.0
    BlocksToSimulate: int = 0
    precision: int = 832
    amplification: float = 10.0
    bitsPerSymbol: int = 1
    # statistics
    runs: int = 0
    errors: int = 0
    reportedErrors: int = 0
    biterrors: int = 0
    BLER: float = 0.0
    BER: float = 0.0
    RER: float = 0.0
    time_sum: float = 0.0
    time_min: float = 0.0
    time_max: float = 0.0
    time_mean: float = 0.0
    time_dev: float = 0.0
    encTime: float = 0.0
    blps: float = 0.0
    cbps: float = 0.0
    pbps: float = 0.0
    ebps: float = 0.0
    effectiveRate: float = 0.0

    def copy(self):
        return dataclasses.replace(self)


def default_data_point(cfg):
    """Simulator::getDefaultDataPoint, simulator.cpp:134-162 (K = int(N * rate) in float arithmetic)"""
    dp = DataPoint()
    dp.designSNR = float(f32(cfg["design-snr"]))
    dp.N = int(cfg["blocklength"])
    dp.K = int(f32(dp.N) * f32(cfg["rate"]))
    dp.L = int(cfg["pathlimit"])
    dp.errorDetection = error_detection_id(cfg["error-detection"])
    dp.errorDetectionType = error_detection_type(cfg["error-detection"])
    dp.systematic = not cfg["non-systematic"]
    dp.EbN0 = float(f32(cfg["snr-max"]))
    dp.BlocksToSimulate = int(cfg["workload"]) // dp.N
    dp.precision = int(cfg["precision"])
    dp.amplification = float(f32(cfg["amplification"]))
    return dp


def _linspace_f32(lo, hi, count):
    """min + i * (max - min) / (count - 1) in float, as the configure*Sim loops compute it"""
    lo, hi = f32(lo), f32(hi)
    scale = (hi - lo) / f32(count - 1)
    return [lo + f32(i) * scale for i in range(count)]


def _configure_single(cfg):
    return [default_data_point(cfg)]


def _configure_codelength(cfg):                                   # simulator.cpp:170-191
    jobs, n = [], int(cfg["n-min"])
    while n <= int(cfg["n-max"]):
        job = default_data_point(cfg)
        job.N = n
        job.K = int(f32(n) * f32(cfg["rate"]))
        job.BlocksToSimulate = int(cfg["workload"]) // n
        jobs.append(job)
        n *= 2
    return jobs


def _configure_designsnr(cfg):                                    # simulator.cpp:193-210
    jobs = []
    for d in _linspace_f32(cfg["dsnr-min"], cfg["dsnr-max"], int(cfg["dsnr-count"])):
        job = default_data_point(cfg)
        job.designSNR = float(d)
        jobs.append(job)
    return jobs


def _configure_listlength(cfg):                                   # simulator.cpp:212-227
    jobs, l = [], int(cfg["l-min"])
    if l < 1:
        raise ValueError("l-min must be >= 1")
    while l <= int(cfg["l-max"]):
        job = default_data_point(cfg)
        job.L = l
        jobs.append(job)
        l *= 2
    return jobs


def _configure_depthfirst(cfg):                                   # simulator.cpp:303-321: the trial limit doubles from l-min to l-max
    jobs = _configure_listlength(cfg)
    for job in jobs:
        job.decoderType = "depthfirst"
    return jobs


def _configure_rate(cfg):                                         # simulator.cpp:229-250
    jobs = []
    for r in _linspace_f32(cfg["r-min"], cfg["r-max"], int(cfg["r-count"])):
        job = default_data_point(cfg)
        job.K = int(f32(job.N) * r)
        job.K = (job.K + 7) // 8 * 8                              # "Round up to nearest byte"
        jobs.append(job)
    return jobs


def _configure_amplification(cfg):                                # simulator.cpp:252-269
    jobs = []
    for a in _linspace_f32(cfg["amp-min"], cfg["amp-max"], int(cfg["amp-count"])):
        job = default_data_point(cfg)
        job.amplification = float(a)
        jobs.append(job)
    return jobs


def _amp_float(ebn0):
    """4 * 10^(Eb/N0 / 10): L_ch for float decoding, simulator.cpp:366-379"""
    return float(f32(4.0 * math.pow(10.0, float(ebn0) / 10.0)))


def _configure_compareall(cfg, skipped=None):                     # simulator.cpp:404-491
    template = default_data_point(cfg)
    template.EbN0 = 2.0
    amp = _amp_float(template.EbN0)
    rows = [("Fast-SSC32", "flexible", 32, amp, 1), ("Fast-SSC8", "flexible", 8, None, 1),
            ("SCL32", "flexible", 32, amp, None), ("SCL8", "flexible", 8, None, None),
            ("SCFlip", "depthfirst", None, amp, None), ("Fast-SSCAN", "fastsscan", None, amp, None)]
    base = []
    for name, dtype, prec, a, l in rows:
        job = template.copy()
        job.name, job.decoderType = name, dtype
        if prec is not None:
            job.precision = prec
        if a is not None:
            job.amplification = a
        if l is not None:
            job.L = l
        base.append(job)
    jobs = []
    amp_hi = _amp_float(3.0)
    for job in base:                                              # "Push two versions of the above list"
        jobs.append(job)
        hi = job.copy()
        hi.N, hi.K, hi.EbN0 = 4096, 2048 + 1024, 3.0
        if hi.precision == 32:
            hi.amplification = amp_hi
        jobs.append(hi)
    fixed = template.copy()
    fixed.name, fixed.decoderType, fixed.amplification = "FFSSC", "fixed", amp_hi
    jobs.append(fixed)
    keep = [j for j in jobs if j.decoderType in ("flexible", "depthfirst")]
    if skipped is not None:
        skipped.extend(j.name for j in jobs if j.decoderType not in ("flexible", "depthfirst"))
    return keep


def snr_inflate(jobs, cfg):
    """Simulator::snrInflateJobList + pushJobsInRange (simulator.cpp:359-402): three ranges [snr-min, 0], [0, 2],
    [2, snr-max] with snr-count/4, /2, /4 grid points each; the FIRST point of every range is not simulated (the
    reference's loops start at i = 1). float precision 32 gets the matched LLR scaling 4 * Eb/N0."""
    out = []
    smin, smax, cnt = f32(cfg["snr-min"]), f32(cfg["snr-max"]), int(cfg["snr-count"])
    for job in jobs:
        for lo, hi, c in ((smin, f32(0.0), cnt // 4), (f32(0.0), f32(2.0), cnt // 2), (f32(2.0), smax, cnt // 4)):
            if c < 2:
                continue                                          # unsigned loop from 1 to c-1 is empty
            scale = (hi - lo) / f32(c - 1)
            for i in range(1, c):
                nj = job.copy()
                nj.EbN0 = float(lo + f32(i) * scale)
                if nj.precision == 32:
                    nj.amplification = _amp_float(nj.EbN0)
                out.append(nj)
    return out


def build_job_list(cfg, skipped=None):
    """Simulator::Simulator, simulator.cpp:41-78"""
    st = cfg["simtype"]
    if st == "compareall":
        return _configure_compareall(cfg, skipped)               # "No SNR inflation"
    table = {"single": _configure_single, "codelength": _configure_codelength, "designsnr": _configure_designsnr,
             "listlength": _configure_listlength, "rate": _configure_rate, "amplification": _configure_amplification,
             "depthfirst": _configure_depthfirst}
    if st == "getcode":
        return []
    if st not in table:
        raise NotImplementedError(f"simulation type '{st}' needs a decoder that has no GPU kernel (out of scope)")
    return snr_inflate(table[st](cfg), cfg)


def decoder_kind(job):
    """SimulationWorker::setCoders, simulator.cpp:703-765 (tFlexible and tDepthFirst branches)"""
    if job.decoderType == "depthfirst":
        return capi.PC_DEC_SCFLIP_F32                            # new DepthFirst(N, L, frozen), :715-716
    if job.decoderType != "flexible":
        raise NotImplementedError(f"decoder type '{job.decoderType}' has no GPU kernel (out of scope)")
    if job.L > 1:
        table = {8: capi.PC_DEC_ADAPTIVE_I8, 32: capi.PC_DEC_ADAPTIVE_F32, 832: capi.PC_DEC_ADAPTIVE_MIXED}
    else:
        table = {8: capi.PC_DEC_FASTSSC_I8, 832: capi.PC_DEC_FASTSSC_I8, 32: capi.PC_DEC_FASTSSC_F32}
    if job.precision not in table:
        raise ValueError(f"No decoder present for {job.precision}-bit decoding.")
    return table[job.precision]


def effective_detector(job):
    """SimulationWorker::setErrorDetector, simulator.cpp:766-819: (bits, type) after the K guard; crc8 / crc32 only"""
    bits, typ = job.errorDetection, job.errorDetectionType
    if bits >= job.K:
        bits, typ = 0, "none"
    if typ == "cmac":
        raise NotImplementedError("CMAC detectors are host-side crypto and out of scope of the GPU path")
    if typ != "crc" or bits not in (8, 32):
        return 0, "none"
    return bits, "crc"


def _dist():
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized():
        return dist, dist.get_rank(), dist.get_world_size()
    return None, 0, 1


def run_job(job, seed=0x5EED, batch=1 << 18, job_index=0, device=None):
    """SimulationWorker::run for one job (simulator.cpp:632-672) on the current CUDA device; fills the statistics."""
    import torch

    from . import pypolar
    from .sweep import shard_range
    dev = torch.device("cuda", torch.cuda.current_device()) if device is None else device
    dist, rank, world = _dist()
    frozen = pypolar.frozen_bits(job.N, job.K, job.designSNR, "BB")                  # selectFrozenBits :686-701
    kind = decoder_kind(job)
    dec = pypolar.PolarDecoder(job.N, job.L, frozen, kind)
    bits, typ = effective_detector(job)
    job.errorDetection, job.errorDetectionType = bits, typ
    dec.setErrorDetection(bits, "crc")
    dec.setSystematic(job.systematic)
    is_float = capi.kind_takes_float(kind)
    N, K, kb = job.N, job.K, (job.K + 7) // 8
    lo, hi = shard_range(job.BlocksToSimulate, rank, world)
    nb = max(1, min(int(batch), hi - lo))
    llr = torch.empty((nb, N), dtype=torch.float32 if is_float else torch.int8, device=dev)
    info = torch.empty((nb, kb), dtype=torch.uint8, device=dev)
    out = torch.empty((nb, kb), dtype=torch.uint8, device=dev)
    ok = torch.empty((nb,), dtype=torch.uint8, device=dev)
    counts = torch.zeros((4,), dtype=torch.int64, device=dev)
    st = ctypes.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
    L = lib()

    def block(first, n, count):
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
        ev[0].record()
        check(L.pc_awgn_generate(dec._h, seed, first, float(job.EbN0), float(job.amplification), int(is_float), ptr(llr),
                                 ptr(info), n, st))
        ev[1].record()
        check(L.pc_decode_batch(dec._h, ptr(llr), int(is_float), ptr(out), ptr(ok), n, st))
        ev[2].record()
        if count:
            check(L.pc_count_errors(ptr(info), ptr(out), ptr(ok), K, n, ctypes.c_void_p(counts.data_ptr()), st))
        return ev

    # warm-up blocks are simulated and not counted (simulator.cpp:642-655); distinct Philox range
    warm = min(job.BlocksToSimulate // 8, 1000, nb)
    if warm:
        block((job_index << 40) + (1 << 39), warm, False)
    events, sizes, f = [], [], lo
    while f < hi:
        n = min(nb, hi - f)
        events.append(block((job_index << 40) + f, n, True))
        sizes.append(n)
        f += n
    torch.cuda.synchronize(dev)
    gen_s = [e[0].elapsed_time(e[1]) * 1e-3 for e in events]
    dec_s = [e[1].elapsed_time(e[2]) * 1e-3 for e in events]
    times = torch.tensor([sum(dec_s), sum(gen_s)], dtype=torch.float64, device=dev)
    if dist is not None and world > 1:
        dist.all_reduce(counts, op=dist.ReduceOp.SUM)
        dist.all_reduce(times, op=dist.ReduceOp.MAX)       # ranks run side by side: the job takes as long as the slowest
    runs, fe, be, rep = [int(v) for v in counts.tolist()]
    job.runs, job.errors, job.biterrors, job.reportedErrors = runs, fe, be, rep
    per_block = np.array([t / n for t, n in zip(dec_s, sizes)], dtype=np.float64) if sizes else np.zeros(1)
    finish_job(job, float(times[0]), float(times[1]), per_block)
    return job


def finish_job(job, dec_time, enc_time, per_block):
    """SimulationWorker::calculateStatistics, simulator.cpp:967-985"""
    job.time_sum, job.encTime = dec_time, enc_time
    job.time_min, job.time_max = float(per_block.min()), float(per_block.max())
    job.time_mean, job.time_dev = float(per_block.mean()), float(per_block.std())
    runs = max(job.runs, 1)
    payload = job.runs * (job.K - job.errorDetection)
    job.BLER = job.errors / runs
    job.BER = job.biterrors / (runs * job.K)
    job.RER = job.reportedErrors / runs
    t = dec_time if dec_time > 0 else float("inf")
    job.blps = job.runs / t
    job.cbps = job.runs * job.N / t
    job.pbps = payload / t
    job.ebps = job.runs * job.N / (enc_time if enc_time > 0 else float("inf"))
    job.effectiveRate = (job.runs - job.errors) * (job.K - job.errorDetection) / t


def _g(x):
    """operator<<(ostream, float): %g with 6 significant digits"""
    return "%g" % float(f32(x))


def format_row(job, compare=False):
    """one CSV line of saveResults / saveComparisonResults (simulator.cpp:535-557, 581-603)"""
    rate = lambda v: _g(v) if v > 0.0 else "1e-99"
    cols = [str(job.N), str(job.K), _g(job.designSNR), str(job.errorDetection), str(job.L), _g(job.EbN0),
            str(job.bitsPerSymbol), rate(job.BLER), rate(job.BER), rate(job.RER), str(job.runs), str(job.errors),
            _g(job.time_sum), _g(job.blps), _g(job.cbps), _g(job.pbps), _g(job.effectiveRate), _g(job.ebps),
            _g(job.amplification), str(int(job.time_min * 1e9)), str(int(job.time_max * 1e9)),
            str(int(job.time_mean * 1e9)), str(int(job.time_dev * 1e9))]
    if compare:
        cols.insert(0, '"' + job.name + '"')
    return ",".join(cols)


def save_results(jobs, cfg):
    """saveResults / saveComparisonResults: <output>_<simtype>.csv"""
    compare = cfg["simtype"] == "compareall"
    name = f"{cfg['output']}_{cfg['simtype']}.csv"
    with open(name, "w") as fh:
        fh.write((CSV_HEADER_COMPARE if compare else CSV_HEADER) + "\n")
        for job in jobs:
            fh.write(format_row(job, compare) + "\n")
    return name


def print_code(cfg, file=sys.stdout):
    """Simulator::printCode, simulator.cpp:493-516: frozen-bit mask, blank line, frozen indices"""
    from . import pypolar
    dp = default_data_point(cfg)
    frozen = pypolar.frozen_bits(dp.N, dp.K, dp.designSNR, "BB")
    fs = set(frozen)
    print("".join("1," if i in fs else "0," for i in range(dp.N)), file=file)
    print(file=file)
    print("".join(f"{i}," for i in frozen), file=file)


def make_parser():
    d = DEFAULTS
    p = argparse.ArgumentParser(prog="pcsim", description="Polar code BLER simulation on B200 (pcsim-compatible)")
    p.add_argument("simtype", nargs="?", default=d["simtype"], choices=SIM_TYPES)
    p.add_argument("-w", "--workload", type=int, default=d["workload"], help="bits per simulation run")
    p.add_argument("--snr-min", type=float, default=d["snr-min"])
    p.add_argument("--snr-max", type=float, default=d["snr-max"])
    p.add_argument("--snr-count", type=int, default=d["snr-count"])
    p.add_argument("-d", "--design-snr", type=float, default=d["design-snr"])
    p.add_argument("--dsnr-min", type=float, default=d["dsnr-min"])
    p.add_argument("--dsnr-max", type=float, default=d["dsnr-max"])
    p.add_argument("--dsnr-count", type=int, default=d["dsnr-count"])
    p.add_argument("-n", "--blocklength", type=int, default=d["blocklength"])
    p.add_argument("--n-min", type=int, default=d["n-min"])
    p.add_argument("--n-max", type=int, default=d["n-max"])
    p.add_argument("-r", "--rate", type=float, default=d["rate"])
    p.add_argument("--r-min", type=float, default=d["r-min"])
    p.add_argument("--r-max", type=float, default=d["r-max"])
    p.add_argument("--r-count", type=int, default=d["r-count"])
    p.add_argument("-l", "--pathlimit", type=int, default=d["pathlimit"])
    p.add_argument("--l-min", type=int, default=d["l-min"])
    p.add_argument("--l-max", type=int, default=d["l-max"])
    p.add_argument("-e", "--error-detection", default=d["error-detection"], choices=ERROR_DETECTIONS)
    p.add_argument("-s", "--non-systematic", action="store_true", default=d["non-systematic"])
    p.add_argument("-p", "--precision", type=int, default=d["precision"], choices=[8, 32, 832])
    p.add_argument("-a", "--amplification", type=float, default=d["amplification"])
    p.add_argument("--amp-min", type=float, default=d["amp-min"])
    p.add_argument("--amp-max", type=float, default=d["amp-max"])
    p.add_argument("--amp-count", type=int, default=d["amp-count"])
    p.add_argument("-o", "--output", default=d["output"])
    p.add_argument("-t", "--threads", type=int, default=d["threads"], help="accepted for compatibility; the GPU batch replaces the worker threads")
    p.add_argument("--batch", type=int, default=1 << 18, help="frames per launch (extension)")
    p.add_argument("--seed", type=lambda s: int(s, 0), default=0x5EED, help="Philox seed (extension)")
    return p


def config_from_args(argv=None):
    ns = make_parser().parse_args(argv)
    return {k.replace("_", "-"): v for k, v in vars(ns).items()}


def main(argv=None):
    cfg = config_from_args(argv)
    if cfg["simtype"] == "getcode":
        print_code(cfg)
        return 0
    skipped = []
    jobs = build_job_list(cfg, skipped)
    for name in skipped:
        print(f"pcsim: skipping '{name}' (decoder has no GPU kernel)", file=sys.stderr)
    _, rank, _ = _dist()
    for i, job in enumerate(jobs):
        run_job(job, seed=cfg["seed"], batch=cfg["batch"], job_index=i)
        if rank == 0:                                            # jobEndingOutput, simulator.cpp:1006-1024 (condensed)
            print(f"[{i + 1}/{len(jobs)}] N={job.N} K={job.K} L={job.L} p={job.precision} Eb/N0={_g(job.EbN0)} dB: "
                  f"BLER={_g(job.BLER)} BER={_g(job.BER)} RER={_g(job.RER)} runs={job.runs} "
                  f"{job.pbps / 1e9:.3f} Gbit/s payload", file=sys.stderr)
    if rank == 0:
        print(save_results(jobs, cfg))
    return 0


if __name__ == "__main__":
    sys.exit(main())
