#!/usr/bin/env python
"""formats the 3GPP TS 38.212 Table 5.3.1.2-1 polar sequence (rank, index pairs in a CSV with 8 pairs per row, the layout
of the reference's python/5g_polar_code_reliability_table.csv) as polar-codes_b200/csrc/pc_5g_table.inc:
tools/make_5g_table.py table.csv > polar-codes_b200/csrc/pc_5g_table.inc"""
import csv
import sys

seq = [None] * 1024
for r in csv.reader(open(sys.argv[1])):
    r = [x for x in r if x.strip() != ""]
    for i in range(0, len(r), 2):
        seq[int(r[i])] = int(r[i + 1])
assert sorted(seq) == list(range(1024))
print("// Polar sequence Q_0^{Nmax-1} of 3GPP TS 38.212 (v15+), Table 5.3.1.2-1: bit-channel indices of an Nmax = 1024 polar code in")
print("// ASCENDING order of reliability (entry i is Q_i). Standard data, shipped as data; tools/make_5g_table.py formats it.")
print("static const unsigned short k5gSequence[1024] = {")
for i in range(0, 1024, 16):
    print("    " + ", ".join(str(v) for v in seq[i:i + 16]) + ",")
print("};")
