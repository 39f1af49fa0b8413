"""Round 2: the wide-alphabet path (5..7 letters: 3-bit Occ blocks, base-8 seed keys, level-synchronous engine) at cfg4 scale.
The cfg4 EDS with a fraction of its letters replaced by N (alphabet {#,A,C,G,N,T}), index files from our GPU builder,
device-resident batches of length-32 patterns, count-only and locate; a sample checked against the unmodified reference
(counts and csv rows).  One JSON line per run -> profiles/r02_wide_alphabet.jsonl.
usage: python profiles/scripts/r02_wide_measure.py [mbp] [npat] [frac_n]"""
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
    mbp = float(sys.argv[1]) if len(sys.argv) > 1 else 250.0
    npat = int(sys.argv[2]) if len(sys.argv) > 2 else 2_000_000
    frac_n = float(sys.argv[3]) if len(sys.argv) > 3 else 0.005
    eb = importlib.import_module("eds-bwt_b200")
    synth = importlib.import_module("eds-bwt_b200.synth")
    from oracle import run as R
    col = synth.chromosome_eds(mbp, seed=4)
    rng = np.random.default_rng(99)
    letters = col.letters.copy()
    where = rng.random(len(letters)) < frac_n
    letters[where] = ord("N")
    col.letters = letters
    out = {"what": "wide alphabet {#,A,C,G,N,T} at cfg4 scale", "mbp": mbp, "frac_N": frac_n, "rows": col.n_rows, "strings": col.nstr, "segments": col.n_seg}
    with tempfile.TemporaryDirectory(prefix="edsbwt_wide_") as d:
        base = os.path.join(d, "wide")
        t0 = time.time()
        eb.build_from_collection(col.letters, col.str_off, col.seg_first, base, device=0)
        out["index_files_built_s"] = round(time.time() - t0, 2)
        t0 = time.time()
        idx = eb.Index(base, verify=False)
        out["open_s"] = round(time.time() - t0, 2)
        info = idx.info()
        out.update(sigma=info.sigma, alphabet=bytes(info.alphabet[:info.sigma]).decode(), seed_k=info.seed_k, engine=info.engine,
                   device_gb=round(info.device_bytes / 1e9, 2), seed_intervals=info.seed_intervals)
        pats = synth.sample_patterns(col, npat, 32, seed=45)
        out["patterns"] = npat
        out["patterns_with_N"] = int((pats == ord("N")).any(axis=1).sum())
        data, off = synth.pack_fixed(pats)
        d_data = torch.from_numpy(data).cuda()
        d_off = torch.from_numpy(off.view(np.int64)).cuda()
        for locate in (False, True):
            ms = []
            for it in range(4):
                r = idx.search_device(d_data.data_ptr(), d_off.data_ptr(), npat, int(data.size), 32, locate=locate)
                st = idx.last_stats()
                ms.append(st["total_ms"])
                if it == 3:
                    counts, occ = r.fetch()
                    out["locate" if locate else "count_only"] = {
                        "ms": round(min(ms[1:]), 3), "patterns_per_s": round(npat / (min(ms[1:]) * 1e-3), 1), "kernel_launches": int(st["kernel_launches"]),
                        "steps": int(st["steps"]), "frontier_peak": int(st["frontier_peak"]), "occurrences": int(counts.sum())}
                r.free()
        # parity on a sample: the unmodified reference (0.5 s per pattern at this size), counts and rows
        ns = 24
        pf = os.path.join(d, "sample.txt")
        with open(pf, "wb") as f:
            for row in pats[:ns]:
                f.write(row.tobytes() + b"\n")
        if R.ref_binary("rb1"):
            t0 = time.time()
            c_ref, rows_ref, took, _ = R.run_ref(base, pf, "rb1")
            c_gpu, o_gpu = idx.search([p.tobytes() for p in pats[:ns]], locate=True)
            ok = c_gpu.tolist() == list(c_ref) and eb.occ_rows(o_gpu) == sorted(rows_ref)
            out["parity_vs_reference"] = ("counts+rows ok" if ok else "MISMATCH") + " (%d patterns, %d rows, reference %.1f s)" % (ns, len(rows_ref), time.time() - t0)
        else:
            out["parity_vs_reference"] = "reference binary not built"
        idx.close()
    print(json.dumps(out))


if __name__ == "__main__":
    main()
