#!/usr/bin/env python
"""BASELINE.json configs 3 and 5 on one B200 (device-side timing, CUDA events inside the library; not a bench.py line).

  cfg3  synthetic SARS-CoV-2-like EDS (30 kbp, 5 000 variant sites), 1 M length-32 patterns, EDSBWTsearch (rank LF) vs
        MOVE_EDSBWTSearch (move-table LF); both variants must return identical counts and occurrence sets
  cfg5  pattern length sweep m = 8 .. 256 on the cfg4 index (250 Mbp), --npat patterns per length per GPU, count-only and
        locate (m = 8 < seed depth: thousands of intervals and occurrences per pattern -- run on a 1/256 sample, stated in the line)

usage: python profiles/scripts/configs_sweep.py [--npat 10000000] [--mbp 250] [--steps 3] > gpurun_out/configs_sweep.jsonl
"""
import argparse
import importlib
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), "..", ".."))
sys.path.insert(0, ROOT)


def log(*a):
    print("[sweep]", *a, file=sys.stderr, flush=True)


def run(idx, torch, data, off, plen, locate, lf, steps):
    d_data = torch.from_numpy(data).cuda()
    d_off = torch.from_numpy(off.view(np.int64)).cuda()
    npat = len(off) - 1
    ms, kms, lms, nocc, launches = [], [], [], 0, 0
    counts = None
    for it in range(2 + steps):
        r = idx.search_device(d_data.data_ptr(), d_off.data_ptr(), npat, int(data.size), plen, locate=locate, lf_variant=lf)
        st = idx.last_stats()
        if it == 0:
            counts, occ = r.fetch() if npat <= 2000000 else (None, None)
        nocc = r.nocc
        r.free()
        if it >= 2:
            ms.append(st["total_ms"]); kms.append(st["search_kernel_ms"]); lms.append(st["locate_kernel_ms"]); launches = st["kernel_launches"]
    return {"ms_per_step": float(np.mean(ms)), "search_kernel_ms": float(np.mean(kms)), "locate_kernel_ms": float(np.mean(lms)),
            "patterns_per_s": npat / (np.mean(ms) * 1e-3), "occurrences": int(nocc), "kernel_launches": int(launches),
            "overflow_patterns": int(st["overflow_patterns"]), "levelsync_patterns": int(st["levelsync_patterns"])}, counts, (occ if npat <= 2000000 else None)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--npat", type=int, default=10000000)
    ap.add_argument("--mbp", type=float, default=250.0)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--skip-cfg3", action="store_true")
    ap.add_argument("--skip-cfg5", action="store_true")
    ap.add_argument("--lens", default="8,16,32,64,128,256")
    ap.add_argument("--fused-cap", type=int, default=0)
    ap.add_argument("--cfg4-move", action="store_true", help="also time MOVE_EDSBWTSearch on the cfg4 index (m = 32)")
    args = ap.parse_args()
    import torch
    eb = importlib.import_module("eds-bwt_b200")
    synth = importlib.import_module("eds-bwt_b200.synth")
    work = os.path.join("/dev/shm" if os.path.isdir("/dev/shm") else "/tmp", "edsbwt_bench_%d" % os.getuid())
    os.makedirs(work, exist_ok=True)

    if not args.skip_cfg3:
        col = synth.sarscov2_like_eds(5000, seed=3)
        base = os.path.join(work, "cfg3")
        eb.build_from_collection(col.letters, col.str_off, col.seg_first, base, device=0)
        pats = synth.sample_patterns(col, 1000000, 32, seed=33)
        data, off = synth.pack_fixed(pats)
        t0 = time.time()
        with eb.Index(base, build_move=True, fused_cap=args.fused_cap) as idx:
            info = idx.info()
            log("cfg3 index: n=%d nText=%d nSeg=%d seed_k=%d move runs=%d intervals=%d (open %.2fs)" % (
                info.n, info.n_text, info.n_seg, info.seed_k, info.move_runs, info.move_intervals, time.time() - t0))
            res = {}
            for name, lf in (("EDSBWTsearch", eb.LF_RANK), ("MOVE_EDSBWTSearch", eb.LF_MOVE)):
                r, counts, occ = run(idx, torch, data, off, 32, True, lf, args.steps)
                res[name] = (counts, eb.occ_rows(occ))
                print(json.dumps({"config": "cfg3: SARS-CoV-2-like EDS (n=%d rows, %d strings, %d segments), 1000000 length-32 patterns, locate" % (
                    info.n, info.n_text, info.n_seg), "variant": name, "seed_k": info.seed_k,
                    "move_runs": info.move_runs, "move_intervals": info.move_intervals, **r}), flush=True)
            same = bool(np.array_equal(res["EDSBWTsearch"][0], res["MOVE_EDSBWTSearch"][0]) and res["EDSBWTsearch"][1] == res["MOVE_EDSBWTSearch"][1])
            print(json.dumps({"config": "cfg3", "variants_identical": same, "occurrences": len(res["EDSBWTsearch"][1])}), flush=True)
            del res

    if args.cfg4_move:
        base = os.path.join(work, "cfg4_%g" % args.mbp)
        col = synth.chromosome_eds(args.mbp, seed=4)
        if not os.path.exists(base + ".done"):
            eb.build_from_collection(col.letters, col.str_off, col.seg_first, base, device=0)
            open(base + ".done", "w").close()
        pats = synth.sample_patterns(col, args.npat, 32, seed=44)
        data, off = synth.pack_fixed(pats)
        del pats, col
        t0 = time.time()
        with eb.Index(base, verify=False, build_move=True, fused_cap=args.fused_cap) as idx:
            info = idx.info()
            log("cfg4 index + move table: %.1fs, runs=%d intervals=%d rounds=%d device_bytes=%.2f GB" % (
                time.time() - t0, info.move_runs, info.move_intervals, info.move_rounds, info.device_bytes / 1e9))
            for name, lf in (("EDSBWTsearch", eb.LF_RANK), ("MOVE_EDSBWTSearch", eb.LF_MOVE)):
                r, _, _ = run(idx, torch, data, off, 32, True, lf, args.steps)
                print(json.dumps({"config": "cfg4 (%g Mbp, n=%d), %d length-32 patterns, locate" % (args.mbp, info.n, args.npat), "variant": name,
                                  "seed_k": info.seed_k, "move_runs": info.move_runs, "move_intervals": info.move_intervals,
                                  "device_bytes": info.device_bytes, **r}), flush=True)
        del data, off

    if not args.skip_cfg5:
        base = os.path.join(work, "cfg4_%g" % args.mbp)
        col = synth.chromosome_eds(args.mbp, seed=4)
        if not os.path.exists(base + ".done"):
            eb.build_from_collection(col.letters, col.str_off, col.seg_first, base, device=0)
            open(base + ".done", "w").close()
        with eb.Index(base, verify=False) as idx:
            info = idx.info()
            for m in [int(x) for x in args.lens.split(",")]:
                # m < seed depth: count-only is one table lookup per pattern; every pattern has thousands of occurrences
                # (4^8 keys over 2.8e8 rows), so locate runs on a 1/256 sample to keep the tuple buffer bounded
                npat = args.npat
                pats = synth.sample_patterns(col, npat, m, seed=55 + m)
                data, off = synth.pack_fixed(pats)
                del pats
                r0, _, _ = run(idx, torch, data, off, m, False, eb.LF_RANK, args.steps)
                r0["patterns"] = npat
                rec = {"config": "cfg5: cfg4 index (%g Mbp, n=%d), %d length-%d patterns" % (args.mbp, info.n, npat, m), "pattern_len": m,
                       "seed_k": info.seed_k, "count_only": r0}
                nloc = npat if m >= 16 else max(1, npat // 256)
                if nloc != npat:
                    data, off = data[: nloc * m].copy(), off[: nloc + 1].copy()
                try:
                    r1, _, _ = run(idx, torch, data, off, m, True, eb.LF_RANK, args.steps)
                    r1["patterns"] = nloc
                    rec["locate"] = r1
                except eb.EdsbwtError as e:
                    rec["locate"] = {"error": str(e), "patterns": nloc}
                print(json.dumps(rec), flush=True)
                del data, off
    return 0


if __name__ == "__main__":
    sys.exit(main())
