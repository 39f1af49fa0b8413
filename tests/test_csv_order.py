"""SURVEY 8(f).3 -- csv emission: the reference's ROW ORDER (unsorted diff against the reference's csv passes) and the
device-side decimal formatter.  Oracle: rows_reference_order of the golden fixtures (written by the unmodified reference),
the unmodified reference / the row-order-exact C restatement on random EDSs, oracle/move_ref.py for the MOVE order."""
import importlib
import os
import subprocess

import numpy as np
import pytest

from conftest import golden_names, load_golden
from oracle import eds_index as E

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def eb():
    return importlib.import_module("eds-bwt_b200")


def _index_files(tmp_path, segs, name="idx"):
    ix = E.build_index(segs)
    base = str(tmp_path / name)
    E.write_index(ix, base)
    return ix, base


def _rows_in_order(occ):
    return list(zip((occ["pat"] + 1).tolist(), occ["str"].tolist(), occ["seg"].tolist(), occ["idx_in_seg"].tolist(), occ["off"].tolist()))


@pytest.mark.parametrize("name", golden_names())
@pytest.mark.parametrize("seed_k,locate_mode,engine,cap", [(-1, 0, 0, 0), (2, 1, 0, 4), (4, 0, 1, 0)])
def test_reference_row_order_golden(eb, tmp_path, name, seed_k, locate_mode, engine, cap):
    g = load_golden(name)
    _, base = _index_files(tmp_path, E.parse_eds(g["eds"]))
    want = [tuple(r) for r in g["rows_reference_order"]]
    with eb.Index(base, seed_k=seed_k, locate_mode=locate_mode, engine=engine, fused_cap=cap) as idx:
        idx.set_row_order(True)
        counts, occ = idx.search(g["patterns"], locate=True)
        assert counts.tolist() == g["counts"]
        assert _rows_in_order(occ) == want
        idx.set_chunk(3)                                   # order must survive the chunked pipeline
        counts, occ = idx.search(g["patterns"], locate=True)
        assert _rows_in_order(occ) == want
        idx.set_row_order(False)
        counts, occ = idx.search(g["patterns"], locate=True)
        assert eb.occ_rows(occ) == sorted(want)


@pytest.mark.parametrize("seed,kw,chunk", [
    (61, dict(n_seg=400, max_var=5, max_len=5, p_empty=0.3), 0),
    (62, dict(n_seg=3000, max_var=4, max_len=6, p_empty=0.15, nondeg_len=(1, 40)), 97),
    (63, dict(n_seg=900, max_var=8, max_len=3, p_empty=0.45, nondeg_len=(1, 3)), 1),
])
def test_search_to_csv_bytes_equal_reference(eb, oracle_built, tmp_path, seed, kw, chunk):
    """whole-file byte equality with the csv the reference writes (unmodified reference when built, else the port)"""
    rng = np.random.default_rng(seed)
    segs = E.random_eds(rng, **kw)
    _, base = _index_files(tmp_path, segs)
    pats = E.sample_patterns(rng, segs, 500, [1, 2, 3, 5, 8, 13, 21, 32, 64])
    pf = str(tmp_path / "p.txt")
    E.write_patterns(pf, pats)
    ref_csv = str(tmp_path / "ref.csv")
    if oracle_built.ref_binary("rb1"):
        c_ref, _, _, _ = oracle_built.run_ref(base, pf, "rb1", keep_csv=ref_csv)
    else:
        subprocess.run([oracle_built.PORT, base, pf, "--csv", ref_csv, "--counts", str(tmp_path / "c")], check=True, capture_output=True)
        c_ref = [int(x) for x in open(tmp_path / "c").read().split()]
    ours = str(tmp_path / "ours.csv")
    with eb.Index(base) as idx:
        idx.set_chunk(chunk)
        counts, nrows = idx.search_to_csv(pats, ours)
        st = idx.last_stats()
    assert counts.tolist() == c_ref and nrows == sum(c_ref)
    assert st["kernel_launches"] > 0
    assert open(ours, "rb").read() == open(ref_csv, "rb").read()


def test_device_formatter_wide_numbers(eb, tmp_path):
    """10-digit pattern numbers (pat_base near 2^32) and every digit count in between, against Python's own formatting"""
    torch = pytest.importorskip("torch")
    rng = np.random.default_rng(7)
    segs = E.random_eds(rng, n_seg=600, max_var=4, max_len=6, p_empty=0.2, nondeg_len=(1, 200))
    _, base = _index_files(tmp_path, segs)
    pats = E.sample_patterns(rng, segs, 700, [2, 3, 4, 6])
    data, off = eb.pack_patterns(pats)
    d_bytes = torch.from_numpy(data.copy()).cuda()
    d_off = torch.from_numpy(off.astype(np.int64)).cuda()
    with eb.Index(base) as idx:
        idx.set_row_order(True)
        for pat_base in (0, 9, 99998, 4294960000):
            res = idx.search_device(d_bytes.data_ptr(), d_off.data_ptr(), len(pats), len(data), 6, locate=True, pat_base=pat_base)
            counts, occ = res.fetch()
            path = str(tmp_path / ("w%d.csv" % pat_base))
            nb = res.write_csv(idx, path)
            res.free()
            want = "#Pat\t$_i\tD[i]\tS_j\tS_j[r]\n" + "".join(
                "%d\t%d\t%d\t%d\t%d\n" % (int(o["pat"]) + 1, o["str"], o["seg"], o["idx_in_seg"], o["off"]) for o in occ)
            got = open(path).read()
            assert got == want
            assert nb == len(want) - len("#Pat\t$_i\tD[i]\tS_j\tS_j[r]\n")
            assert len(occ) > 1000


@pytest.mark.parametrize("name", ["readme_test", "cfg2_examplepaper", "empties_chain", "random_seed1"])
def test_move_rows_in_move_reference_order(eb, tmp_path, name):
    """MOVE_EDSBWTSearch prints rows in frontier order (MOVE_EDSBWTSearch.cpp:326-369): oracle/move_ref.py restates it"""
    from oracle import move_ref as M
    g = load_golden(name)
    ix, base = _index_files(tmp_path, E.parse_eds(g["eds"]))
    ms = M.MoveSearch(ix, a=8)
    want = []
    for i, p in enumerate(g["patterns"]):
        want += ms.search(p.encode() if isinstance(p, str) else p, i + 1)[1]
    with eb.Index(base, build_move=True) as idx:
        counts, occ = idx.search(g["patterns"], locate=True, lf_variant=eb.LF_MOVE)
        assert _rows_in_order(occ) == [tuple(r) for r in want]
        out = str(tmp_path / "m.csv")
        counts2, n = idx.search_to_csv(g["patterns"], out, lf_variant=eb.LF_MOVE)
        assert counts2.tolist() == g["counts"]
        with open(out) as f:
            assert f.readline() == "#Pat\t$_i\tD[i]\tS_j\tS_j[r] \n"
            assert [tuple(int(x) for x in line.split("\t")) for line in f] == [tuple(r) for r in want]
