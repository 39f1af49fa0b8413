"""Round 2: where the e2e leg's time goes.  edsbwt_search_into_packed at cfg4 (10 M length-32 patterns as 2-bit words, pinned host
buffers) for a sweep of pipeline chunk sizes: wall ms per call next to the sum of the chunks' device time (edsbwt_last_stats: total,
search kernel, locate kernel) and the PCIe floor measured in the same process."""
import importlib
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), "..", ".."))
sys.path.insert(0, ROOT)


def main():
    import torch
    npat, plen = 10_000_000, 32
    eb = importlib.import_module("eds-bwt_b200")
    synth = importlib.import_module("eds-bwt_b200.synth")
    workdir = os.path.join("/dev/shm", "edsbwt_bench_%d" % os.getuid())
    os.makedirs(workdir, exist_ok=True)
    base = os.path.join(workdir, "cfg4_250")
    col = synth.chromosome_eds(250, seed=4)
    if not os.path.exists(base + ".done"):
        eb.build_from_collection(col.letters, col.str_off, col.seg_first, base, device=0)
        open(base + ".done", "w").close()
    pats = synth.sample_patterns(col, npat, plen, seed=44)
    del col
    data, off = synth.pack_fixed(pats)
    with eb.Index(base, device=0, verify=False) as idx:
        p_data = eb.pinned_empty(data.size, np.uint8); p_data[:] = data
        p_words = eb.pinned_empty(npat, np.uint64); p_inval = eb.pinned_empty((npat + 31) // 32, np.uint32)
        idx.pack_patterns(p_data, plen, nthreads=os.cpu_count() or 1, words=p_words, invalid=p_inval)
        p_counts = eb.pinned_empty(npat, np.uint32)
        p_occ = eb.pinned_empty(npat + 1024, eb.OCC_DTYPE)
        chunks = [int(c) for c in (sys.argv[1].split(",") if len(sys.argv) > 1 else "262144,524288,1048576,2097152,4194304,10000000".split(","))]
        for chunk in chunks:
            idx.set_chunk(chunk)
            wall = dev = kern = loc = 0.0
            launches = 0
            for it in range(2 + 6):
                t1 = time.perf_counter()
                got = idx.search_into_packed(p_words, p_inval, plen, npat, p_counts, p_occ, locate=True)
                dt = (time.perf_counter() - t1) * 1e3
                st = idx.last_stats()
                if it >= 2:
                    wall += dt; dev += st["total_ms"]; kern += st["search_kernel_ms"]; loc += st["locate_kernel_ms"]; launches = st["kernel_launches"]
            d2h = p_counts.nbytes + got * eb.OCC_DTYPE.itemsize
            print(json.dumps({"chunk_patterns": chunk, "wall_ms": round(wall / 6, 3), "patterns_per_s": round(npat / (wall / 6 * 1e-3)),
                              "sum_of_chunk_device_ms": round(dev / 6, 3), "search_kernel_ms": round(kern / 6, 3), "locate_kernel_ms": round(loc / 6, 3),
                              "kernel_launches": int(launches), "h2d_MB": round((p_words.nbytes + p_inval.nbytes) / 1e6, 1), "d2h_MB": round(d2h / 1e6, 1)}), flush=True)


if __name__ == "__main__":
    main()
