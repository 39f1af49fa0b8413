#!/usr/bin/env python
"""File -> csv through the drop-in executable at cfg4 scale ("end-to-end (file -> csv) reported separately").
Writes the 10 M-line pattern file, runs eds-bwt_b200/EDSBWTsearch <index> <patterns> (stdout discarded: 10 M "Found" lines)
and prints one JSON line with wall time, the executable's own `bs took`, and the csv size."""
import importlib
import json
import os
import re
import subprocess
import sys
import time

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), "..", ".."))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402

eb = importlib.import_module("eds-bwt_b200")
synth = importlib.import_module("eds-bwt_b200.synth")
npat = int(sys.argv[1]) if len(sys.argv) > 1 else 10000000
work = os.path.join("/dev/shm", "edsbwt_bench_%d" % os.getuid())
os.makedirs(work, exist_ok=True)
base = os.path.join(work, "cfg4_250")
col = synth.chromosome_eds(250.0, seed=4)
if not os.path.exists(base + ".done"):
    eb.build_from_collection(col.letters, col.str_off, col.seg_first, base, device=0)
    open(base + ".done", "w").close()
pats = synth.sample_patterns(col, npat, 32, seed=44)
del col
pf = os.path.join(work, "patterns_cfg4.txt")
lines = np.empty((npat, 33), dtype=np.uint8)
lines[:, :32] = pats
lines[:, 32] = 10
lines.tofile(pf)
del lines, pats
exe = os.path.join(ROOT, "eds-bwt_b200", "EDSBWTsearch")
out = {}
# round 2: also with the persisted HBM image (EDSBWT_CACHE=1: the first run builds and writes <index>.edsbwt, the second reads it)
for mode, env in (("locate (RECOVERBW=1)", {}), ("count-only (RECOVERBW=0)", {"EDSBWT_RECOVERBW": "0"}),
                  ("locate, EDSBWT_CACHE=1, first run (writes the image)", {"EDSBWT_CACHE": "1"}),
                  ("locate, EDSBWT_CACHE=1, second run (reads the image)", {"EDSBWT_CACHE": "1"})):
    e = dict(os.environ); e.update(env); e["EDSBWT_VERIFY"] = "0"
    t0 = time.time()
    with open(os.devnull, "w") as null, open(os.path.join(work, "cli_stdout.txt"), "w") as so:
        r = subprocess.run([exe, base, pf], stdout=so, stderr=subprocess.PIPE, text=True, env=e)
    wall = time.time() - t0
    tail = open(os.path.join(work, "cli_stdout.txt")).read()[-200:]
    m = re.search(r"bs took:([0-9.eE+-]+)", tail)
    csv = pf + "output.csv"
    rec = {"mode": mode, "rc": r.returncode, "wall_s": round(wall, 2), "bs_took_s": float(m.group(1)) if m else None,
           "patterns": npat, "found_lines": True, "stderr_tail": r.stderr[-120:]}
    if os.path.exists(base + ".edsbwt"):
        rec["image_file_bytes"] = os.path.getsize(base + ".edsbwt")
    if os.path.exists(csv) and "EDSBWT_RECOVERBW" not in env:
        rec["csv_bytes"] = os.path.getsize(csv)
        with open(csv) as f:
            rec["csv_head"] = [f.readline().rstrip("\n"), f.readline().rstrip("\n")]
        os.remove(csv)
    out[mode] = rec
if os.path.exists(base + ".edsbwt"):
    os.remove(base + ".edsbwt")
print(json.dumps({"config": "cfg4, %d length-32 patterns, eds-bwt_b200/EDSBWTsearch <index> <patterns file>" % npat, **out}))
