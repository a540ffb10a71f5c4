#!/usr/bin/env python
"""bench.py -- decoded information Gbit/s of the batched polar decoders (N=1024, K=512) on B200.

BASELINE.json's metric has two halves on the same code: Fast-SSC int8 (the headline line) and CRC-aided SCL L=8.
Contract (driver): `python bench.py --gpus N --steps K --warmup W` prints ONE JSON line on rank 0.
  value        whole-job decoded info Gbit/s of the headline kernel, LLR batches resident in HBM (CUDA events on the
               launching stream, max over ranks)
  e2e          same metric through the host-buffer C-ABI call pc_decode_batch_host (pinned host LLRs -> H2D -> decode
               -> D2H), at the headline batch size, plus the pure pinned-H2D rate of the same buffer (the PCIe wall)
  roofline     algorithmic bytes (N + K/8 + 1 per frame) / kernel time vs the measured HBM peak
  cpu_baseline the reference's own AVX2 decoder (oracle/_ref) on the host cores decoding a D2H slice of the batch
               the GPU decoded, outputs compared frame by frame (`mismatch_frames`)
  scl          the metric's second half measured the same way: CRC-16 SCL L=8 int8 and fp32 kernels with their own
               value / roofline / cpu_baseline (reference SclFipChar / SclAvxFloat, one object per host thread)
`--impl reference` times the reference CPU path alone (all host threads) on a host-made batch of the same recipe and
prints the same line shape, `scl` included, so every ratio can be formed from the two lines.
`--kind scl_i8 | scl_f32` makes that decoder the headline of the line (both arms).
A step = one decode launch over one batch of B frames per GPU; the Fast-SSC batch (B x 1024 B = 1 GiB at 2^20) is far
larger than the 126 MB L2, so no flush between steps is needed ("l2": "inputs larger than L2"); the SCL batches
(64 / 128 MiB) are written over by 1 GiB of other traffic between kinds but not between steps -- their kernel reads
each row once and is compute bound (12 GB/s of algorithmic traffic), so cache state does not move the number.
"""
import argparse
import ctypes
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

N_CODE, K_CODE, DSNR = 1024, 512, 0.0
EBN0_DB, AMP = 2.0, 10.0

# kind -> (metric, workload string, decoder type string, L, CRC bits, float LLRs, default frames per GPU per step)
KINDS = {
    "fastssc_i8": ("decoded info Gbit/s, N=1024 K=512 Fast-SSC int8",
                   "configs[1]: N=1024 K=512 Fast-SSC int8 'char', BB dSNR 0, CRC-8 detector, BPSK/AWGN 2 dB, amp 10",
                   "char", 1, 8, False, 1 << 20),
    "scl_i8": ("decoded info Gbit/s, N=1024 K=512 CRC-SCL L=8 int8",
               "metric's 2nd half: N=1024 K=512 CRC-16 SCL L=8 int8 'char', BB dSNR 0, BPSK/AWGN 2 dB, amp 10",
               "char", 8, 16, False, 1 << 16),
    "scl_f32": ("decoded info Gbit/s, N=1024 K=512 CRC-SCL L=8 fp32",
                "metric's 2nd half: N=1024 K=512 CRC-16 SCL L=8 fp32 'float', BB dSNR 0, BPSK/AWGN 2 dB, LLR = 4 Eb/N0 y",
                "float", 8, 16, True, 1 << 16),  # 6.9 waves of 148 x 64 frames (2^15 = 3.5 waves leaves the fourth round 46 % full)
}


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200")
    ap.add_argument("--kind", default="fastssc_i8", choices=sorted(KINDS))
    ap.add_argument("--frames", type=int, default=0, help="frames per GPU per step (0: the kind's default)")
    ap.add_argument("--e2e-frames", type=int, default=0, help="frames per e2e step (0: same as --frames)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-secondary", action="store_true")
    return ap.parse_args()


def host_threads():
    """(threads used by the CPU legs, physical cores): one decoder object per hardware thread, as pcsim -t does"""
    logical = os.cpu_count() or 1
    try:
        import psutil
        physical = psutil.cpu_count(logical=False) or logical
    except Exception:
        physical = logical
    return logical, physical


class ClockSampler:
    """samples nvidia-smi clocks and throttle reasons DURING the timed region"""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.rows = []
        self.proc = None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "20", "-i", str(index)], stdout=subprocess.PIPE, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append(line.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm, mx, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            f = [x.strip() for x in r.split(",")]
            if len(f) < 6:
                continue
            try:
                sm.append(float(f[0]))
                mx = float(f[1])
            except ValueError:
                continue
            for n, v in zip(names, f[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons),
                "samples": len(sm)}


def amplification(is_float):
    """pcsim scaling (simulator.cpp:370-379): 4 * Eb/N0 (linear) for the float decoders, the fixed 10 for int8"""
    return 4.0 * 10.0 ** (EBN0_DB / 10.0) if is_float else AMP


# ------------------------------------------------------------------ reference (CPU) legs -----------------------
def make_batch_host(kind, B, seed=7):
    """the recipe of the GPU arm's generator, made on the host with the REFERENCE's encoder: random info -> detector
    generate -> systematic encode (ButterflyFipPacked) -> BPSK 0 -> +1 -> AWGN at EBN0_DB -> LLR = amp * y (int8:
    round to nearest even + saturate, the quantiser of CharContainer::insertLlr)"""
    import refshim
    _, _, _, _, crc, is_float, _ = KINDS[kind]
    fz = refshim.frozen_bits(N_CODE, K_CODE, DSNR)
    rng = np.random.default_rng(seed)
    info = rng.integers(0, 256, (B, K_CODE // 8), dtype=np.uint8)
    code, _ = refshim.encode_batch(N_CODE, fz, info, crc_bits=crc, threads=os.cpu_count() or 1)
    bits = np.unpackbits(code, axis=1)  # MSB first, bit i of the code word
    sigma = 1.0 / np.sqrt(2.0 * (K_CODE / N_CODE) * 10.0 ** (EBN0_DB / 10.0))
    y = (1.0 - 2.0 * bits.astype(np.float32)) + np.float32(sigma) * rng.standard_normal((B, N_CODE), dtype=np.float32)
    y *= np.float32(amplification(is_float))
    if is_float:
        return y.astype(np.float32)
    return np.clip(np.rint(y), -128, 127).astype(np.int8)


def ref_rate(kind, llr, threads, repeat=1):
    """info Gbit/s of the reference's own decode_vector loop (oracle/_ref), one decoder object per thread (long-lived,
    the way pcsim and pcbench use them)"""
    import refshim
    _, _, typ, L, crc, _, _ = KINDS[kind]
    fz = refshim.frozen_bits(N_CODE, K_CODE, DSNR)
    out, ok, secs = refshim.decode_batch(N_CODE, L, fz, typ, llr, crc_bits=crc, threads=threads, repeat=repeat,
                                         return_time=True)
    return llr.shape[0] * repeat * K_CODE / secs / 1e9, secs, out, ok


def ref_timed(kind, steps, warmup, threads, step_seconds=1.0):
    """K timed steps of ~step_seconds of all-thread CPU work each on a host-made batch"""
    frames = (1 << 16) if kind == "fastssc_i8" else (1 << 12)
    llr = make_batch_host(kind, frames)
    _, secs, _, _ = ref_rate(kind, llr, threads, 1)
    repeat = max(1, int(step_seconds / max(secs, 1e-3)))
    for _ in range(warmup):
        ref_rate(kind, llr, threads, repeat)
    t_tot = 0.0
    for _ in range(steps):
        _, secs, _, _ = ref_rate(kind, llr, threads, repeat)
        t_tot += secs
    val = frames * repeat * steps * K_CODE / t_tot / 1e9
    return val, t_tot, f"{frames} frames x {repeat} repeats per step, {threads} threads, one decoder object per thread"


def run_reference(args):
    """--impl reference: the reference's CPU implementation on the box's host cores (rank 0 only)"""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import refshim
    if not refshim.available():
        print(json.dumps({"impl": "reference", "unavailable": "oracle/_ref not built on this box"}))
        return
    threads, physical = host_threads()
    metric, workload, _, _, _, is_float, _ = KINDS[args.kind]
    val, t_tot, sample = ref_timed(args.kind, args.steps, args.warmup, threads)
    line = {
        "impl": "reference", "metric": metric, "value": val, "unit": "Gbit/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * t_tot / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32" if is_float else "int8",
        "data": "synthetic", "config": {"workload": workload},
        "cpu_baseline": {"value": val, "unit": "Gbit/s", "cores": threads, "physical_cores": physical,
                         "kind": "reference", "sample": sample},
        "e2e": {"value": val, "unit": "Gbit/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0}
    if not args.no_secondary and args.kind == "fastssc_i8":
        scl = {}
        for k in ("scl_i8", "scl_f32"):
            v, t, s = ref_timed(k, max(3, min(args.steps, 5)), 1, threads, step_seconds=0.7)
            scl[k] = {"value": v, "unit": "Gbit/s", "cores": threads, "kind": "reference", "sample": s,
                      "config": {"workload": KINDS[k][1]}}
        line["scl"] = scl
    print(json.dumps(line))


# ------------------------------------------------------------------ GPU arm ------------------------------------
def main():
    args = parse()
    if args.impl == "reference":
        return run_reference(args)
    import torch
    import torch.distributed as dist
    import polar_b200_loader
    pcb = polar_b200_loader.load()
    capi = pcb.capi
    L_ = capi.lib()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    dev = torch.device("cuda", local)
    kb = K_CODE // 8
    fz = pcb.frozen_bits(N_CODE, K_CODE, DSNR)
    stream = torch.cuda.current_stream(dev)
    sp = ctypes.c_void_p(stream.cuda_stream)
    threads, physical = host_threads()
    peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(peaks_path):
        peak, peak_src = float(json.load(open(peaks_path))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    else:
        peak, peak_src = 6650.0, "fallback (B200_PROFILING.md)"
    counters_path = os.path.join(ROOT, "profiles", "counters.json")  # ncu --set full summaries, per kernel
    counters = json.load(open(counters_path)) if os.path.exists(counters_path) else {}

    def barrier():
        torch.cuda.synchronize(dev)
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    def allmax(x):
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def timed_kind(kind, frames, steps, warmup, sample_clocks):
        """device-resident decode: synthetic batch from the library's own device-side channel generator (random info
        -> detector generate -> systematic encode -> BPSK -> AWGN -> LLR, Philox counter = global frame index, so a
        rank's shard is its frame range), `steps` launches timed with CUDA events on the launching stream"""
        metric, workload, typ, L, crc, is_float, _ = KINDS[kind]
        d = pcb.PolarDecoder(N_CODE, L, fz, typ)
        d.setErrorDetection(crc, "crc")
        x = torch.empty((frames, N_CODE), dtype=torch.float32 if is_float else torch.int8, device=dev)
        tx = torch.empty((frames, kb), dtype=torch.uint8, device=dev)
        o = torch.empty((frames, kb), dtype=torch.uint8, device=dev)
        v = torch.empty((frames,), dtype=torch.uint8, device=dev)
        capi.check(L_.pc_awgn_generate(d._h, ctypes.c_uint64(0x5EED), ctypes.c_uint64(rank * frames),
                                       ctypes.c_float(EBN0_DB), ctypes.c_float(amplification(is_float)), int(is_float),
                                       capi.ptr(x), capi.ptr(tx), frames, sp))

        def one():
            capi.check(L_.pc_decode_batch(d._h, capi.ptr(x), int(is_float), capi.ptr(o), capi.ptr(v), frames, sp))
        for _ in range(max(warmup, 3)):
            one()
        barrier()
        sampler = ClockSampler(local) if (sample_clocks and rank == 0) else None
        l0 = L_.pc_launch_count()
        evs = [torch.cuda.Event(enable_timing=True) for _ in range(steps + 1)]
        evs[0].record(stream)
        for i in range(steps):
            one()
            evs[i + 1].record(stream)
        barrier()
        launches = int(L_.pc_launch_count() - l0)
        total_ms = allmax(evs[0].elapsed_time(evs[-1]))
        per_launch_ms = [evs[i].elapsed_time(evs[i + 1]) for i in range(steps)]
        clocks = sampler.stop() if sampler else None
        # BLER / detector verdicts through the library's own counting kernel; the one collective north_star names
        # (all_reduce(sum) of the error counters over NCCL) runs here
        counts = torch.zeros(4, dtype=torch.int64, device=dev)
        capi.check(L_.pc_count_errors(capi.ptr(tx), capi.ptr(o), capi.ptr(v), K_CODE, frames, capi.ptr(counts), sp))
        if world > 1:
            dist.all_reduce(counts, op=dist.ReduceOp.SUM)
        cf, ce, cb, cr = [int(t) for t in counts.tolist()]
        value = world * frames * steps * K_CODE / (total_ms * 1e-3) / 1e9
        alg = N_CODE * (4 if is_float else 1) + kb + 1
        avg_launch_s = float(np.mean(per_launch_ms)) * 1e-3
        achieved = frames * alg / avg_launch_s / 1e9
        kname = {"fastssc_i8": "k_fastssc_i8", "scl_i8": "k_scl_i8", "scl_f32": "k_scl_f32"}[kind]
        ctr = counters.get(kname, {})
        traffic = ctr.get("dram_bytes_per_frame")
        roof = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                "traffic": traffic * frames if traffic else None, "peak_source": peak_src, "kernel": kname,
                "alg_bytes_per_frame": alg,
                "issue_active_pct": ctr.get("issue_active_pct"), "alu_pipe_pct": ctr.get("alu_pipe_pct"),
                "counters_source": ctr.get("source"),
                "note": "integer-ALU/issue bound (fp32: FMA/ALU), not HBM bound; see DESIGN.md"}
        res = {"metric": metric, "value": value, "unit": "Gbit/s", "steps": steps, "ms_per_step": total_ms / steps,
               "frames_per_gpu_per_step": frames, "config": {"workload": workload}, "roofline": roof,
               "gpu_launches": launches,
               "check": {"frames": cf, "bler": ce / max(cf, 1), "ber": cb / max(cf * K_CODE, 1),
                         "detector_false_rate": cr / max(cf, 1), "ebn0_db": EBN0_DB,
                         "counted_by": "pc_count_errors + all_reduce(sum)"}}
        return res, clocks, (d, x, o, v)

    def cpu_leg(kind, bufs, target_s):
        """the reference decoding a D2H slice of the batch the GPU just decoded: rate + frame-by-frame comparison"""
        import refshim
        if not refshim.available():
            return {"value": None, "unit": "Gbit/s", "cores": 0, "kind": "reference", "sample": "oracle/_ref not built"}
        _, _, typ, L, crc, is_float, _ = KINDS[kind]
        d, x, o, v = bufs
        n = min(x.shape[0], (1 << 16) if kind == "fastssc_i8" else (1 << 12))
        llr = x[:n].cpu().numpy()
        g_out, g_ok = o[:n].cpu().numpy(), v[:n].cpu().numpy()
        _, secs, r_out, r_ok = ref_rate(kind, llr, threads, 1)
        repeat = max(1, int(target_s / max(secs, 1e-3)))
        val, secs, _, _ = ref_rate(kind, llr, threads, repeat)
        if is_float and L > 1:
            # SclAvxFloat keeps its best path metric across decode() calls: the comparison uses
            # a fresh reference object per frame, the timing above the long-lived object a user would keep
            fzr = refshim.frozen_bits(N_CODE, K_CODE, DSNR)
            m = min(n, 1 << 11)
            r_out, r_ok = refshim.decode_batch(N_CODE, L, fzr, typ, llr[:m], crc_bits=crc, threads=threads, fresh=True)
            g_out, g_ok = g_out[:m], g_ok[:m]
        mism = int(((r_out != g_out).any(axis=1) | (r_ok != g_ok)).sum())
        return {"value": val, "unit": "Gbit/s", "cores": threads, "physical_cores": physical, "kind": "reference",
                "sample": f"the first {n} frames of the GPU's batch (D2H copy) x {repeat} repeats, {secs:.1f} s wall, "
                          "one decoder object per thread",
                "compared_frames": int(len(g_ok)), "mismatch_frames": mism}

    kind = args.kind
    B = args.frames or KINDS[kind][6]
    head, clocks, bufs = timed_kind(kind, B, args.steps, args.warmup, True)
    dec, llr, out, ok = bufs
    is_float = KINDS[kind][5]

    # ---- e2e: host buffers through the C-ABI (pinned LLRs, H2D + decode + D2H inside the timed region)
    Be = min(args.e2e_frames or B, B)
    esz = 4 if is_float else 1
    h_llr = torch.empty((Be, N_CODE), dtype=llr.dtype).pin_memory()
    h_llr.copy_(llr[:Be])
    h_out = torch.empty((Be, kb), dtype=torch.uint8).pin_memory()
    h_ok = torch.empty((Be,), dtype=torch.uint8).pin_memory()

    def e2e_step():
        capi.check(L_.pc_decode_batch_host(dec._h, capi.ptr(h_llr), int(is_float), capi.ptr(h_out), capi.ptr(h_ok), Be))

    for _ in range(2):
        e2e_step()
    barrier()
    e2e_steps = max(3, min(args.steps, 10))
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        e2e_step()
    barrier()
    e2e_s = allmax(time.perf_counter() - t0)
    e2e_val = world * Be * e2e_steps * K_CODE / e2e_s / 1e9
    e2e_match = bool(torch.equal(h_out, out[:Be].cpu()))
    # the PCIe wall beside it: the same pinned buffer copied H2D alone, all ranks at once (aggregate = what the host's
    # root complexes deliver to N GPUs together)
    barrier()
    c0, c1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    c0.record(stream)
    for _ in range(3):
        llr[:Be].copy_(h_llr, non_blocking=True)
    c1.record(stream)
    barrier()
    h2d_s = allmax(c0.elapsed_time(c1) * 1e-3)
    h2d_gbs = world * 3 * Be * N_CODE * esz / h2d_s / 1e9
    e2e = {"value": e2e_val, "unit": "Gbit/s", "h2d_bytes_per_step": Be * N_CODE * esz,
           "d2h_bytes_per_step": Be * (kb + 1), "frames_per_step": Be, "steps": e2e_steps,
           "matches_device_path": e2e_match, "h2d_only_GBps_all_ranks": h2d_gbs,
           "h2d_wall_Gbit_s": h2d_gbs * 8 * K_CODE / (N_CODE * esz * 8),
           "note": "h2d_wall = info Gbit/s if the pinned H2D copy were the only cost (measured here, all ranks at once)"}

    # ---- CPU baseline beside it (rank 0, N=1 only)
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        try:
            cpu = cpu_leg(kind, bufs, 6.0)
        except Exception as e:  # the baseline is a report, never a reason to lose the GPU line
            cpu = {"value": None, "unit": "Gbit/s", "cores": 0, "kind": "reference", "sample": f"failed: {e}"}
    del h_llr, bufs, llr, out, ok, dec
    torch.cuda.empty_cache()

    # ---- the metric's second half (when the headline is Fast-SSC): CRC-SCL L=8 int8 and fp32, measured the same way;
    # plus the other decoder kinds of the path as plain rates
    scl, secondary = None, None
    if kind == "fastssc_i8" and not args.no_secondary:
        scl = {}
        for k in ("scl_i8", "scl_f32"):
            r, _, b2 = timed_kind(k, KINDS[k][6], max(20, min(args.steps, 50)), 3, False)
            if rank == 0 and world == 1 and not args.no_cpu_baseline:
                try:
                    r["cpu_baseline"] = cpu_leg(k, b2, 3.0)
                    if r["cpu_baseline"]["value"]:
                        r["vs_cpu_baseline"] = r["value"] / r["cpu_baseline"]["value"]
                except Exception as e:
                    r["cpu_baseline"] = {"value": None, "sample": f"failed: {e}"}
            scl[k] = r
            del b2
            torch.cuda.empty_cache()

        def plain(kind_id, L, crc, frames, is_f, reps=10):
            d = pcb.PolarDecoder(N_CODE, L, fz, kind_id)
            d.setErrorDetection(crc, "crc")
            x = torch.empty((frames, N_CODE), dtype=torch.float32 if is_f else torch.int8, device=dev)
            tx = torch.empty((frames, kb), dtype=torch.uint8, device=dev)
            o = torch.empty((frames, kb), dtype=torch.uint8, device=dev)
            v = torch.empty((frames,), dtype=torch.uint8, device=dev)
            amp = capi.pcsim_amplification(d._kind, EBN0_DB, AMP)
            capi.check(L_.pc_awgn_generate(d._h, ctypes.c_uint64(0x5EED), ctypes.c_uint64(rank * frames),
                                           ctypes.c_float(EBN0_DB), ctypes.c_float(amp), int(is_f), capi.ptr(x),
                                           capi.ptr(tx), frames, sp))

            def one():
                capi.check(L_.pc_decode_batch(d._h, capi.ptr(x), int(is_f), capi.ptr(o), capi.ptr(v), frames, sp))
            for _ in range(3):
                one()
            barrier()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(stream)
            for _ in range(reps):
                one()
            b.record(stream)
            barrier()
            ms = allmax(a.elapsed_time(b))
            return {"value": world * frames * reps * K_CODE / (ms * 1e-3) / 1e9, "unit": "Gbit/s",
                    "frames_per_gpu_per_step": frames, "steps": reps,
                    "bler": float((o != tx).any(dim=1).float().mean().item())}
        secondary = {
            "adaptive_l8_fp32_crc16": plain("mixed", 8, 16, 1 << 18, True),
            # pcsim's L > 1 decoders at precision 8 / 832 (simulator.cpp:721-733): Fast-SSC int8, failures to the list kernel
            "adaptive_l8_int8_crc16": plain(capi.PC_DEC_ADAPTIVE_I8, 8, 16, 1 << 20, False),  # 6.9 waves of the first pass
            "adaptive_l8_int8_fp32_crc16": plain(capi.PC_DEC_ADAPTIVE_MIXED, 8, 16, 1 << 18, True),
            "fastssc_fp32_crc8": plain("float", 1, 8, 1 << 18, True),
        }

    if rank == 0:
        line = {
            "metric": head["metric"], "value": head["value"], "unit": "Gbit/s", "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": head["ms_per_step"], "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32" if is_float else "int8", "data": "synthetic",
            "config": {"workload": head["config"]["workload"], "frames_per_gpu_per_step": B,
                       "l2": "inputs larger than L2 (1 GiB LLR batch per step)" if kind == "fastssc_i8" else
                             "batch re-read every step; kernel is compute bound (see module docstring)",
                       "parallelism": f"dp{world}"},
            "roofline": head["roofline"], "cpu_baseline": cpu, "e2e": e2e,
            "gpu_launches": head["gpu_launches"], "clocks": clocks, "check": head["check"],
            "scl": scl, "secondary": secondary,
        }
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
