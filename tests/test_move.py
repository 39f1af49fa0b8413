"""MOVE_EDSBWTSearch path: (CPU) the restatement oracle/move_ref.py against every golden vector of the unmodified
EDSBWTsearch -- the only pin the reference offers for this variant -- and (GPU) the move-table kernels
against the same vectors, against the rank-LF kernels, and against the invariants of a valid a-balanced move table."""
import importlib
import os
import re
import subprocess

import numpy as np
import pytest

from conftest import golden_names, load_golden
from oracle import eds_index as E, move_ref as M


# ---- CPU: the oracle itself ------------------------------------------------------------------------------------------
@pytest.mark.parametrize("name", golden_names())
@pytest.mark.parametrize("a", [2, 8, None])
def test_move_restatement_matches_golden(name, a):
    g = load_golden(name)
    if name == "random_seed3" and a != 8:
        pytest.skip("largest fixture: one balancing parameter is enough for the pure-Python restatement")
    ix = E.build_index(E.parse_eds(g["eds"]))
    ms = M.MoveSearch(ix, a=a)
    ms.M.check(ms.ebwt, ms.alphabet)
    counts, rows = [], []
    for i, p in enumerate(g["patterns"]):
        c, r = ms.search(p, i + 1)
        counts.append(c)
        rows += r
    assert counts == g["counts"]
    assert sorted(rows) == sorted(g["rows_reference_order"])


def test_runs_and_lf_pairs_sample():
    """sample/test eBWT: every '#' is its own run; LF pairs = pile start + running counts."""
    ebwt = b"TACTTAGT#TT####T#TTT####A"
    heads, syms = M.runs_of(ebwt)
    assert heads == [0, 1, 2, 3, 5, 6, 7, 8, 9, 11, 12, 13, 14, 15, 16, 17, 20, 21, 22, 23, 24]
    assert bytes(syms) == b"TACTAGT#T####T#T####A"
    q = M.lf_pairs(ebwt, heads, syms, b"#ACGT")
    # piles start at # 0, A 10, C 13, G 14, T 15
    assert q[:8] == [15, 10, 13, 16, 11, 14, 18, 0] and q[-1] == 12


def test_balancing_splits_heavy_intervals():
    # a long run maps onto an output interval full of run heads -> must be cut for small a
    ix = E.build_index([[b"ACGT" * 40], [b"A" * 60, b"C"], [b"ACGT" * 10]])
    ebwt = b"".join(ix.piles)
    t2, t8, t0 = M.MoveTable(ebwt, bytes(ix.alphabet), 2), M.MoveTable(ebwt, bytes(ix.alphabet), 8), M.MoveTable(ebwt, bytes(ix.alphabet), None)
    assert t2.r > t0.r and t2.r >= t8.r >= t0.r
    for t in (t2, t8, t0):
        t.check(ebwt, bytes(ix.alphabet))
    # LF through the table == LF through the unbalanced table for every row
    for row in range(len(ebwt)):
        assert t2.move(row, t2.interval_of(row))[0] == t0.move(row, t0.interval_of(row))[0]


# ---- GPU ---------------------------------------------------------------------------------------------------------------
@pytest.fixture(scope="module")
def eb():
    return importlib.import_module("eds-bwt_b200")


def _index_files(tmp_path, segs, name="idx"):
    ix = E.build_index(segs)
    base = str(tmp_path / name)
    E.write_index(ix, base)
    return ix, base


@pytest.mark.gpu
@pytest.mark.parametrize("name", golden_names())
@pytest.mark.parametrize("seed_k,cap,a", [(-1, 0, 0), (1, 0, 2), (3, 4, 3), (2, 5, 8)])
def test_move_search_golden(eb, tmp_path, name, seed_k, cap, a):
    g = load_golden(name)
    _, base = _index_files(tmp_path, E.parse_eds(g["eds"]))
    with eb.Index(base, seed_k=seed_k, fused_cap=cap, move_a=a, build_move=(a == 2)) as idx:
        counts, occ = idx.search(g["patterns"], locate=True, lf_variant=eb.LF_MOVE)
        assert counts.tolist() == g["counts"]
        assert eb.occ_rows(occ) == sorted(g["rows_reference_order"])
        counts0, _ = idx.search(g["patterns"], locate=False, lf_variant=eb.LF_MOVE)
        assert counts0.tolist() == g["counts"]
        info = idx.info()
        assert info.move_intervals >= info.move_runs > 0 and info.move_a == (a or 8)


@pytest.mark.gpu
@pytest.mark.parametrize("case", range(5))
def test_move_table_invariants(eb, tmp_path, case):
    rng = np.random.default_rng(70 + case)
    segs = [
        E.random_eds(rng, n_seg=400, max_var=5, max_len=6, p_empty=0.25),
        [[b"ACGT" * 40], [b"A" * 60, b"C"], [b"ACGT" * 10]],
        [[b"A" * 500], [b"A" * 499, b"", b"A"], [b"AC" * 300, b"CA" * 300], [b"A" * 150 + b"C"]],
        E.random_eds(rng, n_seg=1500, max_var=3, max_len=4, p_empty=0.1, nondeg_len=(20, 200)),
        [[b"AC"], [b"C", b"A"]],
    ][case]
    ix, base = _index_files(tmp_path, segs)
    ebwt = b"".join(ix.piles)
    for a in (2, 8):
        with eb.Index(base, move_a=a) as idx:
            p, q, dest, sym = idx.move_table()
            info = idx.info()
        ref = M.MoveTable(ebwt, bytes(ix.alphabet), a)
        assert info.move_runs == ref.r_runs
        # same object as the oracle's table class, filled with the GPU's columns -> the oracle's own invariant checker
        t = M.MoveTable.__new__(M.MoveTable)
        t.n, t.r, t.a = len(ebwt), len(p), a
        t.p, t.q, t.dest, t.L = p.tolist() + [len(ebwt)], q.tolist(), dest.tolist(), sym.tolist()
        t.check(ebwt, bytes(ix.alphabet))
        if a == 8 and case in (0, 3, 4):
            assert len(p) == ref.r_runs == ref.r       # nothing heavy: the table IS the run decomposition
        if a == 2 and case in (1, 2):
            assert len(p) > ref.r_runs                  # heavy intervals were split


@pytest.mark.gpu
def test_move_equals_rank_random(eb, oracle_built, tmp_path):
    rng = np.random.default_rng(41)
    segs = E.random_eds(rng, n_seg=3000, max_var=5, max_len=6, p_empty=0.2, nondeg_len=(1, 60))
    _, base = _index_files(tmp_path, segs)
    pats = E.sample_patterns(rng, segs, 1500, [1, 2, 3, 5, 8, 13, 21, 32, 50, 64, 100])
    pf = str(tmp_path / "p.txt")
    E.write_patterns(pf, pats)
    c_port, rows_port, _ = oracle_built.run_port(base, pf, locate=True)
    for seed_k, cap, cap2 in [(-1, 0, 0), (4, 4, 0), (4, 4, 6), (2, 0, 0)]:
        with eb.Index(base, seed_k=seed_k, fused_cap=cap, fused_cap2=cap2, move_a=4) as idx:
            counts, occ = idx.search(pats, locate=True, lf_variant=eb.LF_MOVE)
            st = idx.last_stats()
            assert counts.tolist() == c_port
            assert eb.occ_rows(occ) == sorted(rows_port)
            assert np.all(np.diff(occ["pat"].astype(np.int64)) >= 0)
            if cap == 4:
                assert st["overflow_patterns"] > 0
            c2, o2 = idx.search(pats, locate=True, lf_variant=eb.LF_RANK)       # both variants from one handle
            assert c2.tolist() == c_port and eb.occ_rows(o2) == sorted(rows_port)


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["readme_test", "cfg2_examplepaper", "random_seed1"])
def test_cli_move_dropin(eb, tmp_path, name):
    """MOVE_EDSBWTSearch <index> <patterns>: exit status 1, <patterns>output_M_LF.csv with the trailing-space header
    (MOVE_EDSBWTSearch.cpp:55-59), "num occ N" lines, count_found lines."""
    g = load_golden(name)
    _, base = _index_files(tmp_path, E.parse_eds(g["eds"]))
    pf = str(tmp_path / "kmers.txt")
    E.write_patterns(pf, g["patterns"])
    exe = os.path.join(os.path.dirname(eb.__file__), "MOVE_EDSBWTSearch")
    r = subprocess.run([exe, base, pf], capture_output=True, text=True)
    assert r.returncode == 1, r.stderr
    assert [int(x) for x in re.findall(r"num occ (\d+)", r.stdout)] == [c for c in g["counts"] if c]
    assert r.stdout.count("pattern non presente") == sum(1 for c in g["counts"] if not c)
    nfound = sum(1 for c in g["counts"] if c > 0)
    assert "count_found = %d\n" % nfound in r.stderr and "count_not_found = %d\n" % (len(g["counts"]) - nfound) in r.stderr
    with open(pf + "output_M_LF.csv") as f:
        assert f.readline() == "#Pat\t$_i\tD[i]\tS_j\tS_j[r] \n"
        rows = [tuple(int(x) for x in line.split("\t")) for line in f if line.strip()]
    assert sorted(rows) == sorted(g["rows_reference_order"])
