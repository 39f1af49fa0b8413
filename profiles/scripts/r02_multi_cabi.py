#!/usr/bin/env python
"""edsbwt_search_multi (one process, all visible GPUs, NCCL gather overlapped with the chunked search) at cfg4: 10 M length-32
patterns per GPU from pinned host memory, results gathered in the HBM of device 0.  One JSON line per configuration."""
import importlib
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), "..", ".."))
sys.path.insert(0, ROOT)


def main():
    import ctypes as C
    import torch
    eb = importlib.import_module("eds-bwt_b200")
    synth = importlib.import_module("eds-bwt_b200.synth")
    ng = torch.cuda.device_count()
    mbp = float(os.environ.get("EDSBWT_BENCH_MBP", "250"))
    per_gpu = int(os.environ.get("EDSBWT_BENCH_NPAT", "10000000"))
    workdir = os.path.join("/dev/shm", "edsbwt_bench_%d" % os.getuid())
    os.makedirs(workdir, exist_ok=True)
    base = os.path.join(workdir, "cfg4_%g" % mbp)
    col = synth.chromosome_eds(mbp, seed=4)
    if not os.path.exists(base + ".done"):
        eb.build_from_collection(col.letters, col.str_off, col.seg_first, base, device=0)
        open(base + ".done", "w").close()
    pats = synth.sample_patterns(col, per_gpu * ng, 32, seed=44)
    del col
    data = eb.pinned_empty(pats.size, np.uint8); data[:] = pats.reshape(-1)
    del pats
    single = None
    for p2p in (None, "1") if ng > 1 else (None,):
        os.environ.pop("EDSBWT_MULTI_P2P", None)
        if p2p:
            os.environ["EDSBWT_MULTI_P2P"] = p2p
        t0 = time.time()
        with eb.MultiIndex(base, devices=list(range(ng)), verify=False) as mi:
            t_open = time.time() - t0
            for chunks in (1, 2, 4, 8):
                best = None
                for it in range(4):
                    t1 = time.perf_counter()
                    c, o, info = mi.search(data, locate=True, fixed_len=32, chunks_per_gpu=chunks)
                    wall = (time.perf_counter() - t1) * 1e3
                    if best is None or info["wall_ms"] < best["wall_ms"]:
                        best = dict(info)
                        best["python_wall_ms_incl_fetch"] = wall
                if single is None:
                    single = (int(c.astype(np.int64).sum()), len(o))
                print(json.dumps({"gpus": ng, "transport": "peer copies" if p2p else ("nccl" if mi.nccl else "local"), "chunks_per_gpu": chunks,
                                  "patterns": per_gpu * ng, "occurrences": len(o), "same_totals": (int(c.astype(np.int64).sum()), len(o)) == single,
                                  "call_ms": best["wall_ms"], "patterns_per_s": per_gpu * ng / (best["wall_ms"] * 1e-3), "search_ms_max_over_gpus": best["search_ms"],
                                  "gather_ms": best["gather_ms"], "nvlink_bytes": best["gather_bytes"], "open_s": round(t_open, 2)}), flush=True)


if __name__ == "__main__":
    main()
