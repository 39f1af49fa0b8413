#!/usr/bin/env python
"""Round 2 A/B of the hot kernel at cfg4 (one process, index files and patterns built once):
the item kernel (items.cu), packing fused into the kernel or as a separate pass, against the round-1 step-loop kernel (fused.cu).

    python profiles/scripts/r02_items_ab.py [--mbp 250] [--npat 10000000] [--steps 10] [--variants items:10,items:8,step]
Prints one JSON line per variant (device-resident step, CUDA events on the library's stream)."""
import argparse
import importlib
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), "..", ".."))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--mbp", type=float, default=250)
    ap.add_argument("--npat", type=int, default=10000000)
    ap.add_argument("--len", type=int, default=32, dest="plen")
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--variants", default="step,items,items-pack")
    ap.add_argument("--caps", default="13")
    args = ap.parse_args()
    import torch
    eb = importlib.import_module("eds-bwt_b200")
    synth = importlib.import_module("eds-bwt_b200.synth")
    workdir = os.path.join("/dev/shm", "edsbwt_bench_%d" % os.getuid())
    os.makedirs(workdir, exist_ok=True)
    base = os.path.join(workdir, "cfg4_%g" % args.mbp)
    col = synth.chromosome_eds(args.mbp, seed=4)
    if not os.path.exists(base + ".done"):
        eb.build_from_collection(col.letters, col.str_off, col.seg_first, base, device=0)
        open(base + ".done", "w").close()
    pats = synth.sample_patterns(col, args.npat, args.plen, seed=44)
    del col
    data, off = synth.pack_fixed(pats)
    d_data = torch.from_numpy(data).cuda()
    d_off = torch.from_numpy(off.view(np.int64)).cuda()
    ref_counts = None
    for var in args.variants.split(","):
        for cap in [int(c) for c in args.caps.split(",")]:
            name, _, minb = var.partition(":")
            # items: the kernel packs each short pattern as it takes it; items-pack: separate packing pass first (max_len unknown)
            engine = eb.ENGINE_ITEMS if name.startswith("items") else eb.ENGINE_FUSED_STEP
            max_len = 0 if name == "items-pack" else args.plen
            t0 = time.time()
            with eb.Index(base, device=0, verify=False, engine=engine, fused_cap=cap) as idx:
                t_open = time.time() - t0
                tot = kern = loc = 0.0
                for it in range(3 + args.steps):
                    r = idx.search_device(d_data.data_ptr(), d_off.data_ptr(), args.npat, int(data.size), max_len, locate=True)
                    st = idx.last_stats()
                    if it == 0:
                        counts, occ = r.fetch()
                        if ref_counts is None:
                            ref_counts, ref_sum = counts.copy(), (int(occ["str"].astype(np.uint64).sum()), int(occ["off"].astype(np.uint64).sum()), len(occ))
                        same = bool(np.array_equal(counts, ref_counts)) and ref_sum == (int(occ["str"].astype(np.uint64).sum()), int(occ["off"].astype(np.uint64).sum()), len(occ))
                    r.free()
                    if it >= 3:
                        tot += st["total_ms"]; kern += st["search_kernel_ms"]; loc += st["locate_kernel_ms"]
                print(json.dumps({"variant": var, "cap": cap, "step_ms": tot / args.steps, "search_kernel_ms": kern / args.steps, "locate_ms": loc / args.steps,
                                  "patterns_per_s": args.npat / (tot / args.steps * 1e-3), "overflow": st["overflow_patterns"], "launches": st["kernel_launches"],
                                  "same_as_first_variant": same, "open_s": round(t_open, 2)}), flush=True)


if __name__ == "__main__":
    main()
