#!/usr/bin/env python
"""executed warp instructions and stall samples of an .ncu-rep (--import-source on) summed per enclosing source FUNCTION:
tools/ncu_funcs.py x.ncu-rep   (a line belongs to the last function header above it in its file)"""
import csv, io, os, re, subprocess, sys
rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--print-source", "cuda,sass", "--csv"], capture_output=True, text=True).stdout
fpath, per = None, {}
for r in csv.reader(io.StringIO(raw)):
    if len(r) >= 2 and r[0] in ("File Name", "File Path"):
        fpath = r[1]
        continue
    if len(r) < 8 or r[0] in ("", "Line No"):
        continue
    try:
        ln, samples, inst = int(r[0]), int(r[6]), int(r[7])
    except ValueError:
        continue
    per.setdefault(fpath, []).append((ln, samples, inst))
hdr = re.compile(r"^\s*(?:template\s*<[^>]*>\s*)?(?:static\s+)?(?:__device__|__global__|PC_HD|PC_PHD|__host__|inline)[^;]*?\b([A-Za-z_][A-Za-z0-9_]*)\s*\([^;]*$")
agg, tot_i, tot_s = {}, 0, 0
for path, rows in per.items():
    names = {}
    try:
        cur = "?"
        for i, line in enumerate(open(path), 1):
            m = hdr.match(line)
            if m and not line.strip().startswith("//") and m.group(1) not in ("if", "for", "while", "switch", "return"):
                cur = m.group(1)
            names[i] = cur
    except OSError:
        pass
    for ln, s, i in rows:
        k = (os.path.basename(path or "?"), names.get(ln, "?"))
        a = agg.setdefault(k, [0, 0])
        a[0] += s; a[1] += i
        tot_s += s; tot_i += i
print(f"total: {tot_i} warp instructions, {tot_s} samples")
for (f, fn), (s, i) in sorted(agg.items(), key=lambda kv: -kv[1][1])[:45]:
    print(f"{100.0 * i / max(tot_i, 1):5.1f}% inst {100.0 * s / max(tot_s, 1):5.1f}% smp  {i:11d}  {f}:{fn}")
