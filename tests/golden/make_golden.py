"""Generates tests/golden/*.json by running the UNMODIFIED reference (oracle/_ref/EDSBWTsearch_rb1 and
_rb0, compiled from /root/reference by oracle/Makefile) on index files written by oracle/eds_index.py.

Run from the repo root in the build container:  python tests/golden/make_golden.py
(The fixtures were additionally cross-checked, before being committed, against the reference's own
prebuilt ELF with real sdsl and against the brute-force matcher; see DESIGN.md "Oracle pinning".)
"""
import json
import os
import sys
import tempfile

import numpy as np

sys.path.insert(0, os.path.abspath(os.path.join(os.path.dirname(__file__), "..", "..")))
from oracle import eds_index as E, run as R   # noqa: E402

OUT = os.path.dirname(os.path.abspath(__file__))


sample_patterns = E.sample_patterns


def make(name, eds_text, patterns, note):
    segs = E.parse_eds(eds_text)
    ix = E.build_index(segs)
    with tempfile.TemporaryDirectory() as d:
        base = os.path.join(d, "idx")
        E.write_index(ix, base)
        pf = os.path.join(d, "pat.txt")
        E.write_patterns(pf, patterns)
        c1, rows, _, p1 = R.run_ref(base, pf, "rb1")
        c0, _, _, p0 = R.run_ref(base, pf, "rb0")
    assert p1.returncode == 1 and p0.returncode == 1, (p1.returncode, p0.returncode)
    assert c0 == c1 and len(c1) == len(patterns)
    # RECOVERBW=0 stderr: one "Pattern not found, interrupted in position <posSymb>" per pattern whose frontier dies
    # inside the loop (EDSBWTsearch.cpp:714-717); never printed for single-character patterns
    import re
    pos = [int(x) for x in re.findall(r"interrupted in position (\d+)", p0.stderr)]
    who = [i for i, p in enumerate(patterns) if c0[i] == 0 and len(p) >= 2]
    assert len(pos) == len(who), (name, len(pos), len(who))
    interrupted = [0] * len(patterns)
    for i, v in zip(who, pos):
        interrupted[i] = v
    bf = []
    for i, p in enumerate(patterns):
        bf += [(i + 1,) + t for t in E.brute_force(segs, p)]
    assert sorted(bf) == sorted(rows), name
    rec = {
        "name": name, "note": note, "generator": "oracle/_ref/EDSBWTsearch_rb1 (+_rb0 counts), unmodified reference sources",
        "eds": eds_text.decode() if isinstance(eds_text, bytes) else eds_text,
        "patterns": [p.decode() for p in patterns],
        "counts": c1,
        "interrupted": interrupted,
        "rows_reference_order": rows,
        "ebwt": ix.ebwt.decode(), "n": ix.n, "n_text": ix.n_text,
    }
    with open(os.path.join(OUT, name + ".json"), "w") as f:
        json.dump(rec, f, separators=(",", ":"))
    print(name, ix.n, ix.n_text, len(patterns), sum(c1))


if __name__ == "__main__":
    make("readme_test", b"{TT,TA,C}{T}{TT,A,G}{TT,}{ATT}", [b"TATT", b"ACT", b"TTAT"],
         "README.md:144-167 table (SURVEY App. D.1)")
    make("cfg1_kmers", b"{TT,TA,C}{T}{TT,A,G}{TT,}{ATT}", [b"TTAT", b"ATTC", b"CTTA", b"CATT", b"TCA", b"ATC"],
         "sample/test/kmers.txt lines 1-6 (line 7 holds N: out-of-alphabet, App. C.2) (SURVEY App. D.2)")
    make("cfg2_examplepaper", b"{ATTGCT}{CTA,TA,A}{CTACGGACT}{A,}{CTGT}",
         b"TAC CTA ACTG GCTA TACT A CTGT TTGCTCTAC GCTTAC GCTAC".split(),
         "sample/ExamplePaper + extra patterns (SURVEY App. D.3)")
    for seed, kw in [(1, dict(n_seg=60, max_var=4, max_len=5, p_empty=0.2)),
                     (2, dict(n_seg=400, max_var=6, max_len=4, p_empty=0.35, nondeg_len=(1, 4))),
                     (3, dict(n_seg=1500, max_var=4, max_len=6, p_empty=0.15, nondeg_len=(1, 30)))]:
        rng = np.random.default_rng(seed)
        segs = E.random_eds(rng, **kw)
        pats = sample_patterns(rng, segs, 150, [1, 2, 3, 4, 6, 8, 12, 16, 24, 32] if seed < 3 else [3, 4, 6, 8, 12, 16, 24, 32, 48])
        make("random_seed%d" % seed, E.format_eds(segs), pats, "random EDS, numpy default_rng(%d)" % seed)
    # chain of empties + duplicate strings + single-char segments
    make("empties_chain", b"{ACG}{A,}{C,}{G,}{,T}{ACGT}{A,C,G,T}{,AC,ACG}{TTT}",
         [b"A", b"AC", b"ACG", b"GT", b"GACGT", b"CGA", b"CGC", b"CGG", b"CGT", b"CGACGTA", b"GTTT", b"ACGTTT", b"TACGTTT",
          b"CGAT", b"ACGACGT", b"T", b"TT", b"TTT", b"TTTT", b"G"],
         "chains of segments containing the empty string")
    # ---- alphabets wider than {A,C,G,T} (N / IUPAC codes in real assemblies; Parameters.h lists {$,A,C,G,N,T}).
    #      Named wide_*: used by tests/test_wide_alphabet.py only (the 2-bit kernels and the move table do not take them).
    make("wide_kmers_n", b"{TT,TA,C}{T}{TN,A,G}{NT,}{ANT}",
         [b"TTAT", b"ATNT", b"CTTN", b"NANT", b"TNA", b"N", b"NT", b"TANTANT", b"GANT", b"TTT"],
         "the cfg1 EDS with N in three strings; sample/test/kmers.txt line 7 holds N")
    for seed, alpha, kw in [(11, b"ACGNT", dict(n_seg=300, max_var=5, max_len=5, p_empty=0.3, nondeg_len=(1, 12))),
                            (12, b"ACGNRTY", dict(n_seg=900, max_var=4, max_len=6, p_empty=0.15, nondeg_len=(1, 30)))]:
        rng = np.random.default_rng(seed)
        segs = E.random_eds(rng, alphabet=alpha, **kw)
        pats = sample_patterns(rng, segs, 150, [1, 2, 3, 4, 6, 8, 12, 16, 24, 32, 48])
        make("wide_random_seed%d" % seed, E.format_eds(segs), pats, "random EDS over %s, numpy default_rng(%d)" % (alpha.decode(), seed))

