#!/usr/bin/env python3
"""Pattern sets for the search tools, cut from a multiple sequence alignment (host utility).

Same command line and output format as the reference's extract_patterns_from_msa.py (:43-71): every window of
--length letters of every gap-stripped sequence of a FASTA-formatted MSA is a candidate, --num of them are drawn
without replacement and written one per line, the last line without '\\n' (the layout EDSBWTsearch's getline loop
expects, EDSBWTsearch.cpp:207).  Differences: candidates are drawn by (sequence, offset) index instead of
materialising every window (the reference holds len(MSA) * length bytes of substrings), and --seed makes the draw
reproducible (the reference is unseeded).
"""
import argparse
import bisect
import random


def load_sequences(path):
    seqs, cur = [], []
    with open(path) as f:
        for line in f:
            if line.startswith(">"):
                if cur:
                    seqs.append("".join(cur))
                    cur = []
                continue
            cur.append(line.strip().replace("-", ""))
    if cur:
        seqs.append("".join(cur))
    return seqs


def sample_windows(seqs, length, num, rng):
    """num distinct (sequence, start) windows, uniformly over all windows of all sequences, in draw order."""
    ends, total = [], 0
    for s in seqs:
        total += max(0, len(s) - length + 1)
        ends.append(total)
    if num > total:
        raise ValueError("Sample larger than population or is negative")      # random.sample's own message
    out = []
    for g in rng.sample(range(total), num):
        i = bisect.bisect_right(ends, g)
        start = g - (ends[i - 1] if i else 0)
        out.append(seqs[i][start:start + length])
    return out


def main(argv=None):
    ap = argparse.ArgumentParser()
    ap.add_argument("input")
    ap.add_argument("output")
    ap.add_argument("-l", "--length", type=int, required=True)
    ap.add_argument("-n", "--num", type=int, default=90)
    ap.add_argument("--seed", type=int, default=None)
    args = ap.parse_args(argv)
    pats = sample_windows(load_sequences(args.input), args.length, args.num, random.Random(args.seed))
    with open(args.output, "w") as f:
        f.write("\n".join(pats))


if __name__ == "__main__":
    main()
