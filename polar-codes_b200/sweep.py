"""Device-side BLER sweep: the loop of pcsim's SimulationWorker (reference src/simulation/simulator.cpp:632-672) with
every per-block step on the GPU -- random info, detector generate, encode, BPSK, AWGN, LLR scaling/quantising
(pc_awgn_generate), batch decode (pc_decode_batch), error counting (pc_count_errors). Frames are sharded over ranks
by GLOBAL frame index (the Philox counter), so results do not depend on the number of GPUs; the only collective is
one all_reduce(sum) of the int64 counters at the end.

Statistics follow SimulationWorker::calculateStatistics (simulator.cpp:967-985): BLER = frame errors / runs,
BER = bit errors / (runs * K), RER = reported (detector) errors / runs."""
import ctypes

from . import capi
from .capi import check, lib, ptr

CSV_HEADER = "N,K,dSNR,C,L,Eb/N0,BLER,BER,RER,runs,errors,biterrors,reported,kind"


def shard_range(total, rank, world):
    """contiguous frame range [lo, hi) of `rank`; sizes differ by at most one"""
    base, rem = divmod(int(total), int(world))
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def reduce_counts(counts):
    """all_reduce(sum) of the counter tensor when a process group is up (NCCL on GPUs, gloo in CPU tests)"""
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        dist.all_reduce(counts, op=dist.ReduceOp.SUM)
    return counts


def amplification_for(kind, ebn0_db):
    """pcsim's LLR scaling for a PC_DEC_* kind (or one of the type strings 'char' / 'float' / 'mixed')"""
    if isinstance(kind, str):
        k = kind.lower()
        kind = capi.PC_DEC_FASTSSC_I8 if "char" in k else capi.PC_DEC_SCL_F32
    return capi.pcsim_amplification(kind, ebn0_db)


def run_sweep(decoder, ebn0_points_db, frames_per_point, seed=0x5EED, batch=1 << 18, rank=0, world=1, device=None,
              amplification=None):
    """decoder: pypolar.PolarDecoder. Returns a list of dicts, one per Eb/N0 point (identical on every rank)."""
    import torch
    dev = torch.device("cuda", torch.cuda.current_device()) if device is None else device
    N, K = decoder.blockLength(), decoder.infoLength()
    kb = (K + 7) // 8
    is_float = capi.kind_takes_float(decoder._kind)
    lo, hi = shard_range(frames_per_point, rank, world)
    nb = max(1, min(batch, hi - lo))
    llr = torch.empty((nb, N), dtype=torch.float32 if is_float else torch.int8, device=dev)
    info = torch.empty((nb, kb), dtype=torch.uint8, device=dev)
    out = torch.empty((nb, kb), dtype=torch.uint8, device=dev)
    ok = torch.empty((nb,), dtype=torch.uint8, device=dev)
    counts = torch.zeros((len(ebn0_points_db), 4), dtype=torch.int64, device=dev)
    st = ctypes.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
    L = lib()
    for p, ebn0 in enumerate(ebn0_points_db):
        amp = amplification if amplification is not None else capi.pcsim_amplification(decoder._kind, ebn0)
        f = lo
        while f < hi:
            n = min(nb, hi - f)
            # distinct Philox streams per SNR point: the point index goes into the high bits of the frame counter
            first = (p << 40) + f
            check(L.pc_awgn_generate(decoder._h, seed, first, float(ebn0), float(amp), int(is_float), ptr(llr), ptr(info), n, st))
            check(L.pc_decode_batch(decoder._h, ptr(llr), int(is_float), ptr(out), ptr(ok), n, st))
            check(L.pc_count_errors(ptr(info), ptr(out), ptr(ok), K, n, ctypes.c_void_p(counts[p].data_ptr()), st))
            f += n
    torch.cuda.synchronize(dev)
    counts = reduce_counts(counts)
    rows = []
    for p, ebn0 in enumerate(ebn0_points_db):
        runs, fe, be, rep = [int(v) for v in counts[p].tolist()]
        rows.append({"N": N, "K": K, "L": decoder.listSize(), "EbN0": float(ebn0), "runs": runs, "errors": fe,
                     "biterrors": be, "reported": rep, "BLER": fe / max(runs, 1), "BER": be / max(runs * K, 1),
                     "RER": rep / max(runs, 1)})
    return rows


def to_csv(rows, design_snr=0.0, crc="CRC-8", kind="char"):
    lines = [CSV_HEADER]
    for r in rows:
        lines.append(f"{r['N']},{r['K']},{design_snr},{crc},{r['L']},{r['EbN0']},{r['BLER']:.6e},{r['BER']:.6e},"
                     f"{r['RER']:.6e},{r['runs']},{r['errors']},{r['biterrors']},{r['reported']},{kind}")
    return "\n".join(lines)
