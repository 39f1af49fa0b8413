#!/usr/bin/env python
"""MOVE_EDSBWTSearch on the workload it is for: EDSs of the same size and decreasing run count.
For each backbone (copies x unit, 0.1 % divergence between copies, a variant site every 100 bp): r / n of the eBWT, the size of
the move table, and rank-variant vs move-variant patterns/s for 1 M length-32 patterns, with parity between the two.
    python profiles/scripts/r02_move_repetitive.py > profiles/r02_move_repetitive.jsonl"""
import importlib
import json
import os
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), "..", ".."))
sys.path.insert(0, ROOT)


def main():
    import torch
    eb = importlib.import_module("eds-bwt_b200")
    synth = importlib.import_module("eds-bwt_b200.synth")
    total = int(os.environ.get("EDSBWT_REP_BASES", "30000000"))
    npat = 1000000
    for copies in (1, 10, 100, 1000):
        unit = total // copies
        col = synth.repetitive_eds(unit, copies, 0.001, 100, seed=7)
        with tempfile.TemporaryDirectory(dir="/dev/shm" if os.path.isdir("/dev/shm") else None) as d:
            base = os.path.join(d, "rep")
            t0 = time.time()
            eb.build_from_collection(col.letters, col.str_off, col.seg_first, base)
            t_build = time.time() - t0
            pats = synth.sample_patterns(col, npat, 32, seed=88)
            data, off = synth.pack_fixed(pats)
            d_data = torch.from_numpy(data).cuda(); d_off = torch.from_numpy(off.view(np.int64)).cuda()
            with eb.Index(base, verify=False, build_move=True) as idx:
                info = idx.info()
                res = {}
                for name, lf in (("rank", eb.LF_RANK), ("move", eb.LF_MOVE)):
                    ms = 0.0
                    for it in range(3 + 5):
                        r = idx.search_device(d_data.data_ptr(), d_off.data_ptr(), npat, int(data.size), 32, locate=True, lf_variant=lf)
                        st = idx.last_stats()
                        if it == 0:
                            c, o = r.fetch()
                            res[name] = (c.copy(), np.sort(o, order=["pat", "str", "off"]))
                        r.free()
                        if it >= 3:
                            ms += st["total_ms"]
                    res[name + "_ms"] = ms / 5
                same = bool(np.array_equal(res["rank"][0], res["move"][0]) and np.array_equal(res["rank"][1], res["move"][1]))
                t0 = time.time(); idx.save_move_table(); t_save = time.time() - t0
                mlf_bytes = os.path.getsize(base + "_MLF.edsbwt")
            t0 = time.time()
            with eb.Index(base, verify=False, build_move=True) as idx2:
                t_open_loaded = time.time() - t0
                loaded = bool(idx2.info().from_file & 2)
            print(json.dumps({"copies": copies, "unit_len": unit, "n": int(info.n), "strings": int(info.n_text), "runs": int(info.move_runs),
                              "r_over_n": info.move_runs / info.n, "move_intervals": int(info.move_intervals), "balancing_rounds": int(info.move_rounds),
                              "move_table_file_bytes": mlf_bytes, "index_bytes_with_move": int(info.device_bytes),
                              "rank_ms": res["rank_ms"], "move_ms": res["move_ms"], "rank_patterns_per_s": npat / (res["rank_ms"] * 1e-3),
                              "move_patterns_per_s": npat / (res["move_ms"] * 1e-3), "identical_counts_and_rows": same,
                              "occurrences": int(len(res["rank"][1])), "index_build_s": round(t_build, 2), "mlf_save_s": round(t_save, 2),
                              "open_with_mlf_file_s": round(t_open_loaded, 2), "mlf_loaded": loaded}), flush=True)


if __name__ == "__main__":
    main()
