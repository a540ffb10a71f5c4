#!/usr/bin/env python
"""Golden frozen sets of the 5G construction: generated in the build container from the reference's own table
(python/5g_polar_code_reliability_table.csv) with the filter rule of python/polar_5g_parameters.py:100-125
(get_polar_5g_positions: the sequence restricted to indices < N; that module itself imports an unavailable package,
so its 12-line rule is applied here to its CSV), and checked against the reference's C++ FiveGList at N = 1024.
usage: python tests/golden/make_golden_5g.py   (needs /root/reference and oracle/_ref)"""
import csv
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import refshim  # noqa: E402

seq = [None] * 1024
for r in csv.reader(open("/root/reference/python/5g_polar_code_reliability_table.csv")):
    r = [x for x in r if x.strip() != ""]
    for i in range(0, len(r), 2):
        seq[int(r[i])] = int(r[i + 1])
gold = {}
for N in (32, 64, 128, 256, 512, 1024):
    sub = [v for v in seq if v < N]
    assert sorted(sub) == list(range(N))
    for K in (N // 8, N // 4, N // 2, (3 * N) // 4, N - 1, 1):
        fz = sorted(sub[:N - K])
        if N == 1024:
            assert fz == list(refshim.frozen_bits(N, K, 0.0, "5G"))
        gold[f"{N}_{K}"] = np.array(fz, dtype=np.uint16)
np.savez_compressed(os.path.join(HERE, "fiveg_frozen_sets.npz"), **gold)
print(len(gold), "sets written")
