#!/usr/bin/env python
"""times the batch puncturer kernels (HBM-bound gathers): tools/perf_puncturer.py [N M K frames]; algorithmic bytes =
bytes read + bytes written per code word, against MEASURED_PEAKS.json hbm_gbs"""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import polar_b200_loader
pcb = polar_b200_loader.load()
N, M, K, B = (int(v) for v in (sys.argv[1:5] if len(sys.argv) >= 5 else (1024, 800, 400, 1 << 20)))
peak = 6549.1
try:
    peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]
except Exception:
    pass
p = pcb.Puncturer(M, pcb.frozen_bits(N, K, 0.0))


def timed(fn, reps=20):
    for _ in range(3):
        fn()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / reps


for name, dt, es in (("int8 LLR", torch.int8, 1), ("fp32 LLR", torch.float32, 4)):
    Bn = B if es == 1 else B // 4
    full = torch.randint(-100, 100, (Bn, N), device="cuda").to(dt)
    ms = timed(lambda: p.puncture_batch(full))
    by = Bn * (N + M) * es
    print(f"puncture   {name} N={N}->M={M} B={Bn}: {ms:.3f} ms, {by / ms / 1e6:.0f} GB/s algorithmic ({by / ms / 1e6 / peak:.2f} of {peak:.0f}), {Bn / ms / 1e3:.1f} Mwords/s")
    short = p.puncture_batch(full)
    ms = timed(lambda: p.depuncture_batch(short))
    print(f"depuncture {name} M={M}->N={N} B={Bn}: {ms:.3f} ms, {by / ms / 1e6:.0f} GB/s algorithmic ({by / ms / 1e6 / peak:.2f} of {peak:.0f}), {Bn / ms / 1e3:.1f} Mwords/s")
bits = torch.randint(0, 256, (B, N // 8), device="cuda", dtype=torch.uint8)
ms = timed(lambda: p.puncturePacked_batch(bits))
by = B * (N + M) // 8
print(f"puncturePacked N={N}->M={M} B={B}: {ms:.3f} ms, {by / ms / 1e6:.0f} GB/s algorithmic ({by / ms / 1e6 / peak:.2f} of {peak:.0f}), {B * M / ms / 1e6:.0f} Gbit/s coded")
