#!/usr/bin/env python
"""per-source-line warp-stall samples of an .ncu-rep captured with --import-source on:
tools/ncu_lines.py x.ncu-rep [top]   (ncu -i x --page source --print-source cuda,sass --csv, CUDA rows only)"""
import csv, io, subprocess, sys
rep = sys.argv[1]
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--print-source", "cuda,sass", "--csv"], capture_output=True, text=True).stdout
fname, rows, total = "?", [], 0
for r in csv.reader(io.StringIO(raw)):
    if len(r) >= 2 and r[0] in ("File Name", "File Path"):
        fname = r[1].split("/")[-1]
        continue
    if len(r) < 8 or r[0] in ("", "Line No"):
        continue
    try:
        ln, samples, inst = int(r[0]), int(r[6]), int(r[7])
    except ValueError:
        continue
    if samples or inst:
        rows.append((samples, inst, fname, ln, r[1].strip()))
        total += samples
rows.sort(reverse=True)
print(f"total samples {total}")
for s, i, f, ln, src in rows[:top]:
    print(f"{100.0 * s / total:5.1f}% {s:7d} smp {i:11d} inst  {f}:{ln}  {src[:110]}")
