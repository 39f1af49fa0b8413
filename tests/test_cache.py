"""Persisted tables (csrc/cache.cu): <base>_MLF.edsbwt -- the move table, our counterpart of the reference's
<base>_MLF.aux (build_MLF.cpp:153-161, MOVE_EDSBWTSearch.cpp:186-192) -- and the opt-in image of the whole handle <base>.edsbwt.
A file is used only for the index files it was written for; anything stale or damaged is ignored and the tables are rebuilt."""
import importlib
import os
import shutil
import subprocess

import numpy as np
import pytest

from oracle import eds_index as E

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def eb():
    return importlib.import_module("eds-bwt_b200")


def _case(tmp_path, name="idx", seed=11):
    rng = np.random.default_rng(seed)
    segs = E.random_eds(rng, n_seg=2000, max_var=5, max_len=6, p_empty=0.2, nondeg_len=(1, 60))
    ix = E.build_index(segs)
    d = tmp_path / name
    d.mkdir()
    base = str(d / "x")
    E.write_index(ix, base)
    pats = E.sample_patterns(rng, segs, 1500, [1, 4, 9, 16, 32, 40, 70])
    return base, pats


def test_image_cache_roundtrip_and_staleness(eb, tmp_path):
    base, pats = _case(tmp_path)
    with eb.Index(base) as idx:
        c_ref, o_ref = idx.search(pats, locate=True)
        cm_ref, om_ref = idx.search(pats, locate=True, lf_variant=eb.LF_MOVE)
        k = idx.info().seed_k
    assert not os.path.exists(base + ".edsbwt")
    with eb.Index(base, cache=True, build_move=True) as idx:          # first open: builds, then writes the file
        assert idx.info().from_file == 0
        c, o = idx.search(pats, locate=True)
        assert c.tolist() == c_ref.tolist() and eb.occ_rows(o) == eb.occ_rows(o_ref)
    assert os.path.getsize(base + ".edsbwt") > 4096
    for engine in (eb.ENGINE_FUSED, eb.ENGINE_LEVELSYNC, eb.ENGINE_ITEMS):
        with eb.Index(base, cache=True, engine=engine) as idx:         # later opens: everything comes from the file
            info = idx.info()
            assert info.from_file == 3 and info.seed_k == k and info.move_intervals > 0
            c, o = idx.search(pats, locate=True)
            assert c.tolist() == c_ref.tolist() and eb.occ_rows(o) == eb.occ_rows(o_ref)
            c, o = idx.search(pats, locate=True, lf_variant=eb.LF_MOVE)
            assert c.tolist() == cm_ref.tolist() and eb.occ_rows(o) == eb.occ_rows(om_ref)
            cnt, pos = idx.search_trace(pats)
            assert cnt.tolist() == c_ref.tolist()
    # other options than the file was written with: not used (and, cache being on, rewritten for the new options)
    with eb.Index(base, cache=True, seed_k=2) as idx:
        assert idx.info().from_file == 0 and idx.info().seed_k == 2
        c, o = idx.search(pats, locate=True)
        assert c.tolist() == c_ref.tolist() and eb.occ_rows(o) == eb.occ_rows(o_ref)
    with eb.Index(base, cache=True, seed_k=2) as idx:
        assert idx.info().from_file & 1
    # stale: a source file changed after the cache was written (here: its modification time) -> rebuilt, not trusted
    st = os.stat(base + "_bwt_1.aux")
    os.utime(base + "_bwt_1.aux", ns=(st.st_atime_ns, st.st_mtime_ns + 5_000_000_000))
    with eb.Index(base, cache=True, seed_k=2) as idx:
        assert idx.info().from_file == 0
        c, o = idx.search(pats, locate=True)
        assert c.tolist() == c_ref.tolist()
    # damaged: truncated file -> ignored
    with eb.Index(base, cache=True, seed_k=2) as idx:
        assert idx.info().from_file & 1
    size = os.path.getsize(base + ".edsbwt")
    with open(base + ".edsbwt", "r+b") as f:
        f.truncate(size // 2)
    with eb.Index(base, cache=True, seed_k=2) as idx:
        assert idx.info().from_file == 0
        c, o = idx.search(pats, locate=True)
        assert c.tolist() == c_ref.tolist() and eb.occ_rows(o) == eb.occ_rows(o_ref)
    # cache off: the file is never read
    with eb.Index(base, seed_k=2) as idx:
        assert idx.info().from_file == 0


def test_move_table_file_and_build_mlf(eb, tmp_path):
    base, pats = _case(tmp_path, "mv", seed=12)
    with eb.Index(base, build_move=True, move_a=4) as idx:
        cm_ref, om_ref = idx.search(pats, locate=True, lf_variant=eb.LF_MOVE)
        t_ref = idx.move_table()
        assert idx.info().from_file == 0
    exe = os.path.join(os.path.dirname(eb.__file__), "build_MLF")
    r = subprocess.run([exe, base, "4"], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    assert "M_LF written to " + base + "_MLF.edsbwt" in r.stdout and os.path.exists(base + "_MLF.edsbwt")
    with eb.Index(base) as idx:                          # the first MOVE search loads the table instead of building it
        c, o = idx.search(pats, locate=True, lf_variant=eb.LF_MOVE)
        info = idx.info()
        assert info.from_file == 2 and info.move_a == 4
        assert c.tolist() == cm_ref.tolist() and eb.occ_rows(o) == eb.occ_rows(om_ref)
        assert [o["pat"].tolist(), o["str"].tolist(), o["off"].tolist()] == [om_ref["pat"].tolist(), om_ref["str"].tolist(), om_ref["off"].tolist()]   # MOVE row order too
        t = idx.move_table()
        assert all(np.array_equal(a, b) for a, b in zip(t, t_ref))
    with eb.Index(base, build_move=True, move_a=16) as idx:      # another balancing parameter was asked for: built, not loaded
        assert idx.info().from_file == 0 and idx.info().move_a == 16
    # the drop-in executable finds the table as well; a stale table (index files changed) is rebuilt silently
    pf = str(tmp_path / "p.txt")
    E.write_patterns(pf, pats[:200])
    mexe = os.path.join(os.path.dirname(eb.__file__), "MOVE_EDSBWTSearch")
    r1 = subprocess.run([mexe, base, pf], capture_output=True, text=True)
    csv1 = open(pf + "output_M_LF.csv", "rb").read()
    st = os.stat(base + "_info.aux")
    os.utime(base + "_info.aux", ns=(st.st_atime_ns, st.st_mtime_ns + 7_000_000_000))
    with eb.Index(base) as idx:
        idx.search(pats[:10], locate=True, lf_variant=eb.LF_MOVE)
        assert idx.info().from_file == 0
    r2 = subprocess.run([mexe, base, pf], capture_output=True, text=True)
    assert r1.returncode == 1 and r2.returncode == 1 and open(pf + "output_M_LF.csv", "rb").read() == csv1
    # usage error like the reference's build_MLF
    u = subprocess.run([exe, base], capture_output=True, text=True)
    assert u.returncode == 1 and u.stderr.startswith("usage: ")
