#!/usr/bin/env python
"""Random-access roofline sweep: python profiles/gather_sweep.py > profiles/r02_gather_sweep.txt
One thread per 64-byte block, the search kernels' own 256-bit loads; loads in flight per thread x occupancy x
independent / dependent chains, over (a) a buffer of the size of the cfg4 Occ array, (b) the footprint of the whole cfg4
image, (c) the search kernel's own mix: block fetches over the Occ-sized buffer + 16-byte gathers over 5 GB."""
import importlib
import os
import sys

sys.path.insert(0, os.path.abspath(os.path.join(os.path.dirname(__file__), "..")))
eb = importlib.import_module("eds-bwt_b200")

OCC = (283537175 // 128 + 1) * 64          # 141.8 MB
IMAGE = 20700 << 20
N = 1 << 27
print("# buffer              chains  ilp ctas/SM      GB/s   G req/s")
best = {}
for name, buf, mix_b, mix_f in [("occ-sized 142MB", OCC, 0, 0.0), ("image 20.7GB", IMAGE, 0, 0.0), ("4GiB", 4 << 30, 0, 0.0),
                                ("mix occ+5GB/18%", OCC, 5 << 30, 0.18)]:
    for dep in (False, True):
        for ilp in (1, 2, 4, 8):
            for ctas in (0, 4, 2):
                g, r = eb.gather_bench_cfg(0, buf, N, 3, 64, ilp, ctas, dep, mix_b, mix_f)
                key = (name, dep)
                best[key] = max(best.get(key, 0.0), g)
                print("%-20s %-6s %4d %7s %9.1f %9.2f" % (name, "dep" if dep else "indep", ilp, ctas or "max", g, r / 1e9), flush=True)
print("# best per buffer / chain kind")
for (name, dep), g in best.items():
    print("# %-20s %-6s %9.1f GB/s" % (name, "dep" if dep else "indep", g))
