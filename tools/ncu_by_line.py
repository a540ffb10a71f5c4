#!/usr/bin/env python
"""Joins an ncu SASS source page (--page source --csv) with nvdisasm -gi inline line info of the same kernel and
aggregates executed instructions / stall samples per source line at a chosen inline depth.
usage: ncu_by_line.py <ncu-rep> <cubin> <mangled-kernel-substring> [depth-file-substring]"""
import csv, re, subprocess, sys, collections, io
rep, cubin, kern = sys.argv[1:4]
outer_file = sys.argv[4] if len(sys.argv) > 4 else None
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
hdr = rows[hi]
sass = [dict(zip(hdr, r)) for r in rows[hi + 1:] if len(r) == len(hdr)]
dis = subprocess.run(["nvdisasm", "-gi", "-c", cubin], capture_output=True, text=True).stdout.splitlines()
start = next(i for i, l in enumerate(dis) if l.startswith("\t.section\t.text.") and kern in l)
chain, instrs, cur = [], [], []
for l in dis[start + 1:]:
    if l.startswith("\t.section") or l.startswith("//-----"):
        break
    m = re.match(r'\s*//## File "([^"]+)", line (\d+)', l)
    if m:
        cur.append((m.group(1).split("/")[-1], int(m.group(2))))
        continue
    m = re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(.*?);", l)
    if m:
        if cur:
            chain = cur
            cur = []
        instrs.append((m.group(2).strip(), list(chain)))
assert len(instrs) == len(sass), (len(instrs), len(sass))
inner = collections.Counter(); outer = collections.Counter(); stall = collections.Counter()
tot = 0
for (txt, ch), d in zip(instrs, sass):
    n = int(d["Instructions Executed"]); s = int(d["# Samples"] or 0)
    tot += n
    if not ch:
        continue
    inner[ch[0]] += n
    o = ch[-2] if len(ch) >= 2 else ch[-1]
    if outer_file:
        c = [x for x in ch if outer_file in x[0]]
        o = c[-1] if c else o
    outer[o] += n; stall[o] += s
print("total warp instructions", tot)
print("-- by line in outermost inlined frame (instr %, stall-sample %)")
ts = sum(stall.values()) or 1
for k, v in outer.most_common(40):
    print(f"{v/tot*100:6.2f}% {stall[k]/ts*100:6.2f}%  {k[0]}:{k[1]}")
print("-- by innermost line")
for k, v in inner.most_common(30):
    print(f"{v/tot*100:6.2f}%  {k[0]}:{k[1]}")
