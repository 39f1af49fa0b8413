#!/usr/bin/env python
"""Index open at cfg4: built from the reference's files vs read back from the persisted image (<base>.edsbwt, page cache warm),
and the move table built vs read from <base>_MLF.edsbwt.  One JSON line."""
import importlib
import json
import os
import sys
import time

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), "..", ".."))
sys.path.insert(0, ROOT)
eb = importlib.import_module("eds-bwt_b200")
synth = importlib.import_module("eds-bwt_b200.synth")
work = os.path.join("/dev/shm", "edsbwt_bench_%d" % os.getuid())
os.makedirs(work, exist_ok=True)
base = os.path.join(work, "cfg4_250")
if not os.path.exists(base + ".done"):
    col = synth.chromosome_eds(250.0, seed=4)
    eb.build_from_collection(col.letters, col.str_off, col.seg_first, base, device=0)
    open(base + ".done", "w").close()
    del col
for f in (base + ".edsbwt", base + "_MLF.edsbwt"):
    if os.path.exists(f):
        os.remove(f)
out = {}


def timed(label, **kw):
    t0 = time.time()
    with eb.Index(base, **kw) as idx:
        dt = time.time() - t0
        info = idx.info()
        out[label] = {"open_s": round(dt, 3), "from_file": int(info.from_file), "device_GB": round(info.device_bytes / 1e9, 2)}
        return idx, info


timed("build, verify=1 (the executables' default)", verify=True)
timed("build, verify=0", verify=False)
timed("build + write image (cache=1, first open)", verify=False, cache=True)
out["image_file_GB"] = round(os.path.getsize(base + ".edsbwt") / 1e9, 2)
timed("read image (cache=1, second open)", verify=False, cache=True)
timed("read image (cache=1, third open)", verify=False, cache=True)
os.remove(base + ".edsbwt")
t0 = time.time()
with eb.Index(base, verify=False, build_move=True) as idx:
    out["build incl. move table"] = {"open_s": round(time.time() - t0, 3), "move_intervals": int(idx.info().move_intervals), "device_GB": round(idx.info().device_bytes / 1e9, 2)}
    t1 = time.time(); idx.save_move_table(); out["save move table"] = {"s": round(time.time() - t1, 3), "file_GB": round(os.path.getsize(base + "_MLF.edsbwt") / 1e9, 2)}
t0 = time.time()
with eb.Index(base, verify=False, build_move=True) as idx:
    out["open with <base>_MLF.edsbwt"] = {"open_s": round(time.time() - t0, 3), "from_file": int(idx.info().from_file)}
os.remove(base + "_MLF.edsbwt")
print(json.dumps(out))
