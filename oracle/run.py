"""TEST INFRASTRUCTURE (oracle) -- runners for the CPU checkers.

  run_port(...)  oracle/edsbwt_oracle      our C restatement (always available once built)
  run_ref(...)   oracle/_ref/EDSBWTsearch_rb{0,1} / _prebuilt   the unmodified reference (when built)

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may call these.
"""
import json
import os
import re
import shutil
import subprocess
import tempfile

HERE = os.path.dirname(os.path.abspath(__file__))
PORT = os.path.join(HERE, "edsbwt_oracle")
REF_DIR = os.path.join(HERE, "_ref")


def ref_binary(kind="rb1"):
    p = os.path.join(REF_DIR, "EDSBWTsearch_" + kind)
    return p if os.path.exists(p) else None


def build(quiet=True):
    """make port (+ ref when /root/reference is present)."""
    subprocess.run(["make", "-C", HERE, "all"], check=True,
                   stdout=subprocess.DEVNULL if quiet else None, stderr=subprocess.STDOUT if quiet else None)


def run_port(base, patterns_file, locate=True, threads=1, k=12, limit=None):
    """-> (counts list, sorted rows list or None, stats dict)."""
    with tempfile.TemporaryDirectory() as d:
        cmd = [PORT, base, patterns_file, "--counts", d + "/c", "--locate", "1" if locate else "0",
               "--threads", str(threads), "--k", str(k), "--interrupted", d + "/i"]
        if locate:
            cmd += ["--csv", d + "/o.csv"]
        if limit is not None:
            cmd += ["--limit", str(limit)]
        r = subprocess.run(cmd, capture_output=True, text=True, check=True)
        stats = json.loads(r.stdout.strip().splitlines()[-1])
        counts = [int(x) for x in open(d + "/c").read().split()]
        stats["interrupted"] = [int(x) for x in open(d + "/i").read().split()]
        rows = None
        if locate:
            rows = []
            with open(d + "/o.csv") as f:
                f.readline()
                for line in f:
                    rows.append(tuple(int(x) for x in line.split("\t")))
        return counts, rows, stats


def run_ref(base, patterns_file, kind="rb1", keep_csv=None):
    """Run the unmodified reference. -> (counts, rows in reference order or None, 'bs took' seconds, proc)."""
    exe = ref_binary(kind)
    if exe is None:
        raise FileNotFoundError("oracle/_ref/EDSBWTsearch_%s not built" % kind)
    with tempfile.TemporaryDirectory() as d:
        pf = os.path.join(d, "p.txt")
        shutil.copyfile(patterns_file, pf)
        r = subprocess.run([exe, base, pf], capture_output=True, text=True)
        counts = [int(x) for x in re.findall(r"Found (\d+) starting", r.stdout)]
        m = re.search(r"bs took:([0-9.eE+-]+)", r.stdout)
        took = float(m.group(1)) if m else None
        rows = None
        csv = pf + "output.csv"
        if os.path.exists(csv):
            rows = []
            with open(csv) as f:
                f.readline()
                for line in f:
                    if line.strip():
                        rows.append(tuple(int(x) for x in line.split("\t")))
            if keep_csv:
                shutil.copyfile(csv, keep_csv)
        return counts, rows, took, r
