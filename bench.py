#!/usr/bin/env python
"""bench.py -- decoded information Gbit/s of the batched Fast-SSC int8 decoder (N=1024, K=512), B200.

Contract (driver): `python bench.py --gpus N --steps K --warmup W` prints ONE JSON line on rank 0.
  value        whole-job decoded info Gbit/s, LLR batches resident in HBM (CUDA events, max over ranks)
  e2e          same metric through the host-buffer C-ABI call (pinned host LLRs -> H2D -> decode -> D2H)
  roofline     algorithmic bytes (N + K/8 + 1 per frame) / kernel time vs measured HBM peak
  cpu_baseline the reference's own AVX2 Fast-SSC char decoder (oracle/_ref) on the host cores, bounded sample
`--impl reference` times that CPU path alone (all host threads) and prints the same line shape.
A step = one decode launch over one batch of B frames per GPU; the batch (B x 1024 B = 1 GiB at 2^20) is far
larger than the 126 MB L2, so no flush between steps is needed ("l2": "inputs larger than L2").
"""
import argparse
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
METRIC = "decoded info Gbit/s, N=1024 K=512 Fast-SSC int8"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200")
    ap.add_argument("--frames", type=int, default=1 << 20, help="frames per GPU per step")
    ap.add_argument("--e2e-frames", type=int, default=1 << 18)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-secondary", action="store_true")
    return ap.parse_args()


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


def make_llr_host(rng, B):
    """all-zero codeword over BPSK/AWGN at EBN0_DB, int8 LLR = sat(rint(AMP*y)) (pcsim defaults)"""
    sigma = 1.0 / np.sqrt(2.0 * (K_CODE / N_CODE) * 10.0 ** (EBN0_DB / 10.0))
    y = 1.0 + sigma * rng.standard_normal((B, N_CODE), dtype=np.float32)
    return np.clip(np.rint(AMP * y), -128, 127).astype(np.int8)


def cpu_reference_rate(frames, threads, repeat=1):
    """info Gbit/s of the reference's own decode_vector loop (oracle/_ref), one decoder object per thread"""
    import refshim
    fz = refshim.frozen_bits(N_CODE, K_CODE, DSNR)
    llr = make_llr_host(np.random.default_rng(7), frames)
    _, _, secs = refshim.decode_batch(N_CODE, 1, fz, "char", llr, threads=threads, repeat=repeat, return_time=True)
    return frames * repeat * K_CODE / secs / 1e9, secs


def run_reference(args):
    """--impl reference: the reference's CPU implementation on the box's host cores (rank 0 only)"""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import refshim
    cores = os.cpu_count() or 1
    frames = 1 << 16
    if not refshim.available():
        print(json.dumps({"impl": "reference", "unavailable": "oracle/_ref not built on this box"}))
        return
    # size one step to ~1 s of all-core work
    g, secs = cpu_reference_rate(frames, cores, 1)
    repeat = max(1, int(1.0 / max(secs, 1e-3)))
    for _ in range(args.warmup):
        cpu_reference_rate(frames, cores, repeat)
    t_tot, n_tot = 0.0, 0
    for _ in range(args.steps):
        g, secs = cpu_reference_rate(frames, cores, repeat)
        t_tot += secs
        n_tot += frames * repeat
    val = n_tot * K_CODE / t_tot / 1e9
    sample = f"{frames} frames x {repeat} repeats per step, {cores} threads, one decoder object per thread"
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": val, "unit": "Gbit/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * t_tot / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "int8", "data": "synthetic",
        "config": {"workload": "configs[1]: N=1024 K=512 Fast-SSC int8 'char', BPSK/AWGN 2 dB, amp 10"},
        "cpu_baseline": {"value": val, "unit": "Gbit/s", "cores": cores, "kind": "reference", "sample": sample},
        "e2e": {"value": val, "unit": "Gbit/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0}))


def main():
    args = parse()
    if args.impl == "reference":
        return run_reference(args)
    import torch
    import torch.distributed as dist
    import polar_b200_loader
    pcb = polar_b200_loader.load()
    capi = pcb.capi
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    dev = torch.device("cuda", local)
    B = args.frames
    kb = K_CODE // 8
    fz = pcb.frozen_bits(N_CODE, K_CODE, DSNR)
    dec = pcb.PolarDecoder(N_CODE, 1, fz, "char")
    # synthetic batch from the library's own device-side channel generator (random info -> CRC-8 -> systematic
    # encode -> BPSK -> AWGN -> int8 LLR, Philox counter = global frame index, so a rank's shard is its frame range)
    import ctypes
    llr = torch.empty((B, N_CODE), dtype=torch.int8, device=dev)
    sent = torch.empty((B, kb), dtype=torch.uint8, device=dev)
    capi.check(capi.lib().pc_awgn_generate(dec._h, ctypes.c_uint64(0x5EED), ctypes.c_uint64(rank * B), ctypes.c_float(EBN0_DB),
                                           ctypes.c_float(AMP), 0, capi.ptr(llr), capi.ptr(sent), B, None))
    torch.cuda.synchronize(dev)
    out = torch.empty((B, kb), dtype=torch.uint8, device=dev)
    ok = torch.empty((B,), dtype=torch.uint8, device=dev)
    stream = torch.cuda.current_stream(dev)
    sp = __import__("ctypes").c_void_p(stream.cuda_stream)

    def step():
        capi.check(capi.lib().pc_decode_batch(dec._h, capi.ptr(llr), 0, capi.ptr(out), capi.ptr(ok), B, sp))

    def barrier():
        torch.cuda.synchronize(dev)
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    for _ in range(max(args.warmup, 3)):
        step()
    barrier()
    sampler = ClockSampler(local) if rank == 0 else None
    l0 = capi.lib().pc_launch_count()
    evs = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps + 1)]
    evs[0].record(stream)
    for i in range(args.steps):
        step()
        evs[i + 1].record(stream)
    barrier()
    launches = capi.lib().pc_launch_count() - l0
    total_ms = evs[0].elapsed_time(evs[-1])
    per_launch_ms = [evs[i].elapsed_time(evs[i + 1]) for i in range(args.steps)]
    clocks = sampler.stop() if sampler else None
    t = torch.tensor([total_ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    total_ms_max = float(t.item())
    value = world * B * args.steps * K_CODE / (total_ms_max * 1e-3) / 1e9
    bler = float((out != sent).any(dim=1).float().mean().item())
    verdict_false = float((ok == 0).float().mean().item())

    # ---- secondary: the metric's other half, CRC-aided SCL L=8 (N=1024 K=512, CRC-16), int8 and fp32 kernels,
    # LLRs resident in HBM, whole-job info Gbit/s (max over ranks); plus the fp32 Fast-SSC and the adaptive decoder
    def timed_kind(kind, L, crc_bits, frames, is_float, reps):
        d = pcb.PolarDecoder(N_CODE, L, fz, kind)
        d.setErrorDetection(crc_bits, "crc")
        x = torch.empty((frames, N_CODE), dtype=torch.float32 if is_float else torch.int8, device=dev)
        tx = torch.empty((frames, kb), dtype=torch.uint8, device=dev)
        o = torch.empty((frames, kb), dtype=torch.uint8, device=dev)
        v = torch.empty((frames,), dtype=torch.uint8, device=dev)
        # pcsim scaling: 4 * Eb/N0 for the float decoders, the fixed amplification for int8 and for AdaptiveMixed (-p 832)
        amp = 4.0 * 10.0 ** (EBN0_DB / 10.0) if (is_float and kind != capi.PC_DEC_ADAPTIVE_MIXED) else AMP
        capi.check(capi.lib().pc_awgn_generate(d._h, ctypes.c_uint64(0x5EED), ctypes.c_uint64(rank * frames),
                                               ctypes.c_float(EBN0_DB), ctypes.c_float(amp), int(is_float), capi.ptr(x),
                                               capi.ptr(tx), frames, None))

        def one():
            capi.check(capi.lib().pc_decode_batch(d._h, capi.ptr(x), int(is_float), capi.ptr(o), capi.ptr(v), frames, sp))
        for _ in range(2):
            one()
        barrier()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(stream)
        for _ in range(reps):
            one()
        b.record(stream)
        barrier()
        tt = torch.tensor([a.elapsed_time(b)], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        return {"value": world * frames * reps * K_CODE / (float(tt.item()) * 1e-3) / 1e9, "unit": "Gbit/s",
                "frames_per_gpu_per_step": frames, "steps": reps, "bler": float((o != tx).any(dim=1).float().mean().item())}

    secondary = None
    if not args.no_secondary:
        secondary = {
            "scl_l8_int8_crc16": timed_kind("char", 8, 16, 1 << 16, False, 5),
            "scl_l8_fp32_crc16": timed_kind("float", 8, 16, 1 << 15, True, 5),
            "adaptive_l8_fp32_crc16": timed_kind("mixed", 8, 16, 1 << 17, True, 5),
            # pcsim's L > 1 decoders at precision 8 / 832 (simulator.cpp:721-733): Fast-SSC int8, failures to the list kernel
            "adaptive_l8_int8_crc16": timed_kind(capi.PC_DEC_ADAPTIVE_I8, 8, 16, 1 << 18, False, 5),
            "adaptive_l8_int8_fp32_crc16": timed_kind(capi.PC_DEC_ADAPTIVE_MIXED, 8, 16, 1 << 17, True, 5),
            "fastssc_fp32_crc8": timed_kind("float", 1, 8, 1 << 17, True, 5),
        }

    # ---- e2e: host buffers through the C-ABI (pinned LLRs, H2D + decode + D2H inside the timed region)
    Be = min(args.e2e_frames, B)
    h_llr = torch.empty((Be, N_CODE), dtype=torch.int8).pin_memory()
    h_llr.copy_(llr[:Be].cpu())
    h_out = torch.empty((Be, kb), dtype=torch.uint8).pin_memory()
    h_ok = torch.empty((Be,), dtype=torch.uint8).pin_memory()

    def e2e_step():
        capi.check(capi.lib().pc_decode_batch_host(dec._h, capi.ptr(h_llr), 0, capi.ptr(h_out), capi.ptr(h_ok), Be))

    for _ in range(2):
        e2e_step()
    barrier()
    e2e_steps = max(3, min(args.steps, 10))
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        e2e_step()
    barrier()
    e2e_s = time.perf_counter() - t0
    te = torch.tensor([e2e_s], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(te, op=dist.ReduceOp.MAX)
    e2e_val = world * Be * e2e_steps * K_CODE / float(te.item()) / 1e9
    e2e_match = bool(torch.equal(h_out, out[:Be].cpu()))

    # ---- CPU baseline beside it (rank 0, N=1 only): the reference's AVX2 decoder on the host cores
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        try:
            import refshim
            if refshim.available():
                cores = os.cpu_count() or 1
                frames = 1 << 16
                _, secs = cpu_reference_rate(frames, cores, 1)
                repeat = max(1, int(6.0 / max(secs, 1e-3)))
                val, secs = cpu_reference_rate(frames, cores, repeat)
                one, _ = cpu_reference_rate(frames, 1, max(1, repeat // cores))
                cpu = {"value": val, "unit": "Gbit/s", "cores": cores, "kind": "reference",
                       "sample": f"{frames} frames x {repeat} repeats, {secs:.1f} s wall, one decoder per thread",
                       "per_core": one}
        except Exception as e:  # the baseline is a report, never a reason to lose the GPU line
            cpu = {"value": None, "unit": "Gbit/s", "cores": 0, "kind": "reference", "sample": f"failed: {e}"}

    if rank == 0:
        peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
        if os.path.exists(peaks_path):
            peak, peak_src = float(json.load(open(peaks_path))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        else:
            peak, peak_src = 6650.0, "fallback (B200_PROFILING.md)"
        alg_bytes = B * (N_CODE + kb + 1)
        avg_launch_s = float(np.mean(per_launch_ms)) * 1e-3
        achieved = alg_bytes / avg_launch_s / 1e9
        traffic = None
        tpath = os.path.join(ROOT, "profiles", "traffic.json")
        if os.path.exists(tpath):
            tj = json.load(open(tpath))
            if tj.get("frames_per_launch"):
                traffic = tj["dram_bytes_per_launch"] * (B / tj["frames_per_launch"])
        line = {
            "metric": METRIC, "value": value, "unit": "Gbit/s", "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": total_ms_max / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "int8", "data": "synthetic",
            "config": {"workload": "configs[1]: N=1024 K=512 Fast-SSC int8 'char', BB dSNR 0, CRC-8 detector, "
                                   "BPSK/AWGN 2 dB, amp 10", "frames_per_gpu_per_step": B,
                       "l2": "inputs larger than L2 (1 GiB LLR batch per step)", "parallelism": f"dp{world}"},
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": traffic, "peak_source": peak_src, "kernel": "k_fastssc_i8",
                         "alg_bytes_per_frame": N_CODE + kb + 1, "note": "integer-ALU/issue bound, see DESIGN.md"},
            "cpu_baseline": cpu,
            "e2e": {"value": e2e_val, "unit": "Gbit/s", "h2d_bytes_per_step": Be * N_CODE,
                    "d2h_bytes_per_step": Be * (kb + 1), "frames_per_step": Be, "matches_device_path": e2e_match},
            "gpu_launches": int(launches), "clocks": clocks,
            "check": {"bler": bler, "detector_false_rate": verdict_false, "ebn0_db": EBN0_DB},
            "secondary": secondary,
        }
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
