#!/usr/bin/env python
"""Random-gather throughput of this GPU by granule and working-set size (roofline denominators of DESIGN.md section 8).
    python profiles/gather_sweep.py > profiles/rNN_gather_sweep.txt       (needs the GPU)"""
import importlib
import os
import sys

sys.path.insert(0, os.path.abspath(os.path.join(os.path.dirname(__file__), "..")))
eb = importlib.import_module("eds-bwt_b200")
print("# random gather, GB/s (algorithmic bytes = accesses x granule), best of 3, 2^26 accesses")
print("%-12s %10s %10s %10s" % ("buffer", "32 B", "64 B", "128 B"))
for mb in [64, 142, 512, 4096, 16384, 65536]:
    row = []
    for g in (32, 64, 128):
        row.append(eb.gather_bench(0, mb << 20, 1 << 26, 3, g))
    print("%-12s %10.0f %10.0f %10.0f" % ("%d MiB" % mb, *row))
