#!/usr/bin/env python
"""per-launch times of one decoder configuration (CUDA events around every launch), to see run-to-run spread:
tools/launch_times.py N K L kind crc frames ebn0 launches"""
import ctypes, sys, os, subprocess, threading, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import polar_b200_loader
pcb = polar_b200_loader.load()
N, K, L, kind, crc, B, ebn0, R = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3]), sys.argv[4], int(sys.argv[5]), int(sys.argv[6]), float(sys.argv[7]), int(sys.argv[8])
fz = pcb.frozen_bits(N, K, 0.0)
dec = pcb.PolarDecoder(N, L, fz, kind)
dec.setErrorDetection(crc, "crc")
isf = "char" not in kind
llr = torch.empty((B, N), dtype=torch.float32 if isf else torch.int8, device="cuda")
info = torch.empty((B, (K + 7) // 8), dtype=torch.uint8, device="cuda")
amp = 4.0 * 10 ** (ebn0 / 10) if isf else 10.0
pcb.capi.check(pcb.capi.lib().pc_awgn_generate(dec._h, 1, 0, ctypes.c_float(ebn0), ctypes.c_float(amp), int(isf), pcb.capi.ptr(llr), pcb.capi.ptr(info), B, None))
samples = []
stop = False
def poll():
    while not stop:
        try:
            o = subprocess.run(["nvidia-smi", "--query-gpu=clocks.sm,power.draw,clocks_throttle_reasons.active", "--format=csv,noheader"], capture_output=True, text=True, timeout=5).stdout.strip()
            samples.append(o)
        except Exception:
            pass
        time.sleep(0.05)
th = threading.Thread(target=poll); th.start()
out, ok = dec.decode_batch(llr, return_ok=True); torch.cuda.synchronize()
ev = [torch.cuda.Event(enable_timing=True) for _ in range(R + 1)]
ev[0].record()
for i in range(R):
    out, ok = dec.decode_batch(llr, return_ok=True)
    ev[i + 1].record()
torch.cuda.synchronize()
stop = True; th.join()
ms = [ev[i].elapsed_time(ev[i + 1]) for i in range(R)]
print("per-launch ms:", " ".join(f"{x:.2f}" for x in ms))
print(f"min {min(ms):.3f} median {sorted(ms)[R // 2]:.3f} max {max(ms):.3f} mean {sum(ms) / R:.3f} -> {B * K / (sum(ms) / R) / 1e6:.1f} Gbit/s")
print("clock/power samples:", sorted(set(samples))[:12])
