"""GPU parity tests: the CUDA path (through the C ABI, via ctypes) against the oracle on the same inputs.
Bit-exact bar: identical counts, identical occurrence sets, identical byte-model counters."""
import importlib
import os

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


@pytest.mark.parametrize("name", golden_names())
@pytest.mark.parametrize("seed_k,locate_mode,engine,cap", [(-1, 0, 0, 0), (1, 0, 0, 0), (3, 1, 0, 4), (6, 0, 1, 0), (2, 0, 0, 5),
                                                           (-1, 0, 2, 0), (2, 1, 2, 5), (-1, 0, 3, 0), (1, 0, 3, 0), (3, 1, 3, 4), (2, 0, 3, 5)])
def test_golden(eb, tmp_path, name, seed_k, locate_mode, engine, cap):
    g = load_golden(name)
    _, base = _index_files(tmp_path, E.parse_eds(g["eds"]))
    with eb.Index(base, seed_k=seed_k, locate_mode=locate_mode, engine=engine, fused_cap=cap) as idx:
        counts, occ = idx.search(g["patterns"], locate=True)
        assert counts.tolist() == g["counts"]
        assert eb.occ_rows(occ) == sorted(g["rows_reference_order"])
        assert np.all(np.diff(occ["pat"].astype(np.int64)) >= 0)          # grouped by ascending pattern
        counts0, _ = idx.search(g["patterns"], locate=False)
        assert counts0.tolist() == g["counts"]


def _mixed_patterns(rng, segs, n, lens):
    return E.sample_patterns(rng, segs, n, lens)


@pytest.mark.parametrize("seed,kw,seed_k", [
    (21, dict(n_seg=300, max_var=5, max_len=5, p_empty=0.3), 2),
    (22, dict(n_seg=2000, max_var=4, max_len=6, p_empty=0.15, nondeg_len=(1, 40)), 4),
    (23, dict(n_seg=1200, max_var=8, max_len=3, p_empty=0.45, nondeg_len=(1, 3)), 5),
    (24, dict(n_seg=5000, max_var=3, max_len=8, p_empty=0.1, nondeg_len=(5, 120)), -1),
])
def test_random_vs_port(eb, oracle_built, tmp_path, seed, kw, seed_k):
    rng = np.random.default_rng(seed)
    segs = E.random_eds(rng, **kw)
    _, base = _index_files(tmp_path, segs)
    pats = _mixed_patterns(rng, segs, 600, [1, 2, 3, 5, 8, 13, 21, 32, 50, 64, 100])
    pats += [b"", b"ACGNT", b"#A", b"A" * 300]
    pf = str(tmp_path / "p.txt")
    E.write_patterns(pf, [p if p else b"A" for p in pats])       # the file format cannot hold an empty line
    file_pats = [p if p else b"A" for p in pats]
    c_port = rows_port = None
    # the fused kernel (default), the fused kernel with a tiny frontier capacity (most patterns fall through to the
    # level-synchronous engine), and the level-synchronous engine alone: same counts, rows and byte-model counters
    for engine, cap, cap2 in [(eb.ENGINE_FUSED, 0, 0), (eb.ENGINE_FUSED, 4, 0), (eb.ENGINE_FUSED, 4, 7), (eb.ENGINE_LEVELSYNC, 0, 0),
                              (eb.ENGINE_FUSED_STEP, 0, 0), (eb.ENGINE_FUSED_STEP, 4, 7), (eb.ENGINE_ITEMS, 0, 0), (eb.ENGINE_ITEMS, 4, 0), (eb.ENGINE_ITEMS, 4, 7)]:
        with eb.Index(base, seed_k=seed_k, locate_mode=0, engine=engine, fused_cap=cap, fused_cap2=cap2) as idx:
            k = idx.info().seed_k
            idx.set_stats(True, canonical_k=k)
            if c_port is None:
                c_port, rows_port, st_port = oracle_built.run_port(base, pf, locate=True, k=k)
            counts, occ = idx.search(file_pats, locate=True)
            st = idx.last_stats()
            assert counts.tolist() == c_port
            assert eb.occ_rows(occ) == sorted(rows_port)
            assert np.all(np.diff(occ["pat"].astype(np.int64)) >= 0)
            for key in "ISELO":
                assert st[key] == st_port[key], (engine, cap, key, st[key], st_port[key])
            if engine != eb.ENGINE_LEVELSYNC and cap == 4:
                assert st["overflow_patterns"] > 0
                assert cap2 == 0 or st["levelsync_patterns"] > 0
            idx.set_stats(False)
            counts, occ = idx.search(file_pats, locate=True)
            assert counts.tolist() == c_port and eb.occ_rows(occ) == sorted(rows_port)
            # the empty pattern itself (not representable in a file) -> 0
            c2, o2 = idx.search([b"", b"ACG"], locate=True)
            assert c2[0] == 0 and all(o2["pat"] == 1)
    with eb.Index(base, seed_k=seed_k, locate_mode=1) as idx:          # LF-walk locate
        counts, occ = idx.search(file_pats, locate=True)
        assert counts.tolist() == c_port and eb.occ_rows(occ) == sorted(rows_port)


def test_edge_cases(eb, tmp_path):
    # single segment (the reference aborts here, App. C.3): plain matching
    _, base = _index_files(tmp_path, [[b"ACGTACGT", b"CGT"]], "one")
    with eb.Index(base) as idx:
        c, o = idx.search([b"CGT", b"T", b"TA", b"GG"])
        assert c.tolist() == [3, 3, 1, 0]
        assert eb.occ_rows(o) == sorted([(1, 0, 1, 0, 1), (1, 0, 1, 0, 5), (1, 1, 1, 1, 0), (2, 0, 1, 0, 3), (2, 0, 1, 0, 7),
                                         (2, 1, 1, 1, 2), (3, 0, 1, 0, 3)])
        c, o = idx.search([], locate=True)
        assert len(c) == 0 and len(o) == 0
    # fewer than 4 letters
    segs = [[b"AAC"], [b"A", b"", b"CA"], [b"CCA"]]
    _, base = _index_files(tmp_path, segs, "two")
    with eb.Index(base, seed_k=2) as idx:
        pats = [b"A", b"AC", b"CAC", b"ACC", b"ACAC", b"AACCCA", b"G", b"T"]
        c, o = idx.search(pats)
        exp = []
        for i, p in enumerate(pats):
            bf = E.brute_force(segs, p) if all(ch in b"AC" for ch in p) else []
            assert c[i] == len(bf)
            exp += [(i + 1,) + t for t in bf]
        assert eb.occ_rows(o) == sorted(exp)


def test_bad_index_files(eb, tmp_path):
    ix, base = _index_files(tmp_path, E.parse_eds(b"{TT,TA,C}{T}{TT,A,G}{TT,}{ATT}"))
    with open(base + "_bwt_1.aux", "ab") as f:
        f.write(b"A")
    with pytest.raises(eb.EdsbwtError) as ei:
        eb.Index(base)
    assert ei.value.code == -2
    with pytest.raises(eb.EdsbwtError) as ei:
        eb.Index(str(tmp_path / "missing"))
    assert ei.value.code == -1


def test_csv_and_pattern_file(eb, tmp_path):
    g = load_golden("readme_test")
    _, base = _index_files(tmp_path, E.parse_eds(g["eds"]))
    pf = str(tmp_path / "kmers.txt")
    with open(pf, "wb") as f:
        f.write(b"TATT\nACT\nTTAT")            # last line without '\n'
    data, off = eb.read_patterns(pf)
    assert len(off) == 4 and bytes(data[off[2]:off[3]]) == b"TTAT"
    with eb.Index(base) as idx:
        counts, occ = idx.search((data, off))
    eb.write_csv(pf + "output.csv", occ)
    header, rows = E.read_csv_rows(pf + "output.csv")
    assert header == "#Pat\t$_i\tD[i]\tS_j\tS_j[r]\n"
    assert sorted(rows) == sorted(g["rows_reference_order"])


# ---- GPU index builder (edsbwt_build_from_collection / _from_eds): byte-identical files -------------------------
def _files(base):
    out = {}
    for suf in ["_info.aux", ".bitvector"] + ["_bwt_%d.aux" % p for p in range(8)] + ["_bv_%d.aux" % p for p in range(8)]:
        if os.path.exists(base + suf):
            with open(base + suf, "rb") as f:
                out[suf] = f.read()
    return out


def _builder_cases():
    rng = np.random.default_rng(77)
    cases = [E.parse_eds(load_golden(n)["eds"]) for n in golden_names()]
    cases.append(E.random_eds(rng, n_seg=3000, max_var=6, max_len=7, p_empty=0.2, nondeg_len=(1, 60)))
    # long repeats: many doubling rounds, duplicate strings, strings that are prefixes/suffixes of each other
    cases.append([[b"A" * 300], [b"A" * 299, b"A" * 300, b"", b"A"], [b"ACAC" * 80, b"CACA" * 80, b"ACAC" * 80], [b"A" * 150 + b"C"]])
    cases.append([[b"ACGT" * 50, b"ACGT" * 50], [b"T"], [b"ACGT" * 49 + b"ACG", b""], [b"GT" * 100]])
    cases.append([[b"AC"], [b"C", b"A"]])                                  # two letters only
    cases.append([[b"A"], [b"AA", b""], [b"AAA"]])                         # one letter
    return cases


@pytest.mark.parametrize("case", range(len(_builder_cases())))
def test_builder_files_identical_to_oracle_writer(eb, tmp_path, case):
    segs = _builder_cases()[case]
    ix, base = _index_files(tmp_path, segs, "ora")
    synth = importlib.import_module("eds-bwt_b200.synth")
    col = synth.from_segments(segs)
    gbase = str(tmp_path / "gpu")
    eb.build_from_collection(col.letters, col.str_off, col.seg_first, gbase)
    a, b = _files(base), _files(gbase)
    assert sorted(a) == sorted(b)
    for k in a:
        assert a[k] == b[k], k
    # the .eds text route
    eds = str(tmp_path / "x.eds")
    with open(eds, "wb") as f:
        f.write(E.format_eds([[s if s else b"E" for s in seg] if len(seg) == 1 else seg for seg in segs]))
    ebase = str(tmp_path / "eds")
    eb.build_from_eds(eds, ebase)
    assert _files(ebase) == a


def test_synthetic_generators_search_parity(eb, oracle_built, tmp_path):
    """cfg3-shaped and cfg4-shaped workloads at reduced scale: GPU builder -> GPU search vs the oracle."""
    synth = importlib.import_module("eds-bwt_b200.synth")
    for col, npat in [(synth.sarscov2_like_eds(600, seed=3), 3000), (synth.chromosome_eds(0.2, seed=4), 3000)]:
        base = str(tmp_path / col.name.replace(".", "_"))
        eb.build_from_collection(col.letters, col.str_off, col.seg_first, base)
        pats = synth.sample_patterns(col, npat, 32, seed=33)
        pf = base + ".pat"
        with open(pf, "wb") as f:
            for r in pats:
                f.write(r.tobytes() + b"\n")
        c_port, rows_port, _ = oracle_built.run_port(base, pf, locate=True)
        with eb.Index(base) as idx:
            counts, occ = idx.search(synth.pack_fixed(pats), locate=True)
        assert counts.tolist() == c_port
        assert eb.occ_rows(occ) == sorted(rows_port)
        assert sum(c_port) > 0.5 * npat


def test_search_into_pipelined_chunks(eb, oracle_built, tmp_path):
    """Caller-owned (pinned) buffers, batch cut into pipeline chunks: same counts / rows as one call; E_NOSPACE."""
    rng = np.random.default_rng(91)
    segs = E.random_eds(rng, n_seg=1500, max_var=4, max_len=6, p_empty=0.15, nondeg_len=(1, 40))
    _, base = _index_files(tmp_path, segs)
    pats = E.sample_patterns(rng, segs, 1000, [4, 9, 16, 32, 33, 64])
    data, off = eb.pack_patterns(pats)
    with eb.Index(base) as idx:
        c_ref, o_ref = idx.search(pats, locate=True)
        p_data = eb.pinned_empty(data.size, np.uint8); p_data[:] = data
        p_off = eb.pinned_empty(off.size, np.uint64); p_off[:] = off
        for chunk in [0, 1000, 333, 64, 1]:
            idx.set_chunk(chunk)
            counts = eb.pinned_empty(len(pats), np.uint32); counts[:] = 0xffffffff
            occ = eb.pinned_empty(len(o_ref) + 5, eb.OCC_DTYPE)
            n = idx.search_into(p_data, p_off, counts, occ, locate=True)
            assert n == len(o_ref) and counts.tolist() == c_ref.tolist()
            assert eb.occ_rows(occ[:n]) == eb.occ_rows(o_ref)
            assert np.all(np.diff(occ[:n]["pat"].astype(np.int64)) >= 0)
            counts[:] = 0
            assert idx.search_into(p_data, p_off, counts, None, locate=False) == 0 and counts.tolist() == c_ref.tolist()
        idx.set_chunk(100)
        small = eb.pinned_empty(max(1, len(o_ref) // 2), eb.OCC_DTYPE)
        with pytest.raises(eb.EdsbwtError) as ei:
            idx.search_into(p_data, p_off, counts, small, locate=True)
        assert ei.value.code == -7
        # plain (pageable) numpy buffers work too
        idx.set_chunk(1 << 21)
        counts2 = np.zeros(len(pats), np.uint32); occ2 = np.zeros(len(o_ref), eb.OCC_DTYPE)
        assert idx.search_into(data, off, counts2, occ2) == len(o_ref) and eb.occ_rows(occ2) == eb.occ_rows(o_ref)


@pytest.mark.parametrize("plen", [1, 7, 16, 32, 33, 70])
def test_search_into_fixed_length_batches(eb, tmp_path, plen):
    """edsbwt_search_into_fixed (offsets generated on the device) == edsbwt_search_into with an explicit offset array;
    odd lengths exercise the unaligned pattern-window paths of the fused kernel"""
    rng = np.random.default_rng(300 + plen)
    segs = E.random_eds(rng, n_seg=1200, max_var=4, max_len=6, p_empty=0.15, nondeg_len=(1, 60))
    _, base = _index_files(tmp_path, segs)
    pats = E.sample_patterns(rng, segs, 900, [plen])
    assert all(len(p) == plen for p in pats)
    data, off = eb.pack_patterns(pats)
    with eb.Index(base) as idx:
        c_ref, o_ref = idx.search(pats, locate=True)
        exp = []
        for i, p in enumerate(pats[:60]):
            exp.append(len(E.brute_force(segs, p)))
        assert c_ref[:60].tolist() == exp
        for chunk in [0, 257, 1]:
            idx.set_chunk(chunk)
            counts = eb.pinned_empty(len(pats), np.uint32); counts[:] = 0xffffffff
            occ = eb.pinned_empty(len(o_ref) + 3, eb.OCC_DTYPE)
            p_data = eb.pinned_empty(data.size, np.uint8); p_data[:] = data
            n = idx.search_into_fixed(p_data, plen, counts, occ, locate=True)
            assert n == len(o_ref) and counts.tolist() == c_ref.tolist()
            assert eb.occ_rows(occ[:n]) == eb.occ_rows(o_ref)


# ---- patterns handed in as 2-bit words (edsbwt_search_into_packed + the host packer edsbwt_pack_patterns) --------
@pytest.mark.parametrize("plen", [1, 7, 16, 31, 32, 33, 64, 70])
def test_search_into_packed(eb, tmp_path, plen):
    """packed entry point == byte entry points, for every engine tier (the tiers that read characters get them unpacked on
    the device), incl. patterns holding a byte outside the alphabet (flagged by the packer -> 0 occurrences)"""
    rng = np.random.default_rng(500 + plen)
    segs = E.random_eds(rng, n_seg=1200, max_var=4, max_len=6, p_empty=0.15, nondeg_len=(1, 60))
    _, base = _index_files(tmp_path, segs)
    pats = E.sample_patterns(rng, segs, 900, [plen])
    for i in (3, 77, 500, 899):                                   # bytes outside {A,C,G,T}, incl. the end-marker itself
        b = bytearray(pats[i]); b[rng.integers(0, plen)] = [ord("N"), ord("#"), 0, 255][i % 4]; pats[i] = bytes(b)
    data, off = eb.pack_patterns(pats)
    for engine, cap, lf in [(eb.ENGINE_FUSED, 0, eb.LF_RANK), (eb.ENGINE_FUSED, 4, eb.LF_RANK), (eb.ENGINE_LEVELSYNC, 0, eb.LF_RANK),
                            (eb.ENGINE_FUSED_STEP, 0, eb.LF_RANK), (eb.ENGINE_ITEMS, 0, eb.LF_RANK), (eb.ENGINE_FUSED, 0, eb.LF_MOVE)]:
        with eb.Index(base, engine=engine, fused_cap=cap) as idx:
            c_ref, o_ref = idx.search(pats, locate=True, lf_variant=lf)
            assert all(c_ref[i] == 0 for i in (3, 77, 500, 899))
            words, invalid, nbad = idx.pack_patterns(data, plen, nthreads=3)
            assert nbad == 4 and [i for i in range(len(pats)) if (invalid[i >> 5] >> (i & 31)) & 1] == [3, 77, 500, 899]
            p_words = eb.pinned_empty(words.size, np.uint64); p_words[:] = words
            for chunk in [0, 256, 97, 1]:
                idx.set_chunk(chunk)
                counts = eb.pinned_empty(len(pats), np.uint32); counts[:] = 0xffffffff
                occ = eb.pinned_empty(len(o_ref) + 3, eb.OCC_DTYPE)
                n = idx.search_into_packed(p_words, invalid, plen, len(pats), counts, occ, locate=True, lf_variant=lf)
                assert n == len(o_ref) and counts.tolist() == c_ref.tolist(), (engine, cap, lf, chunk)
                assert eb.occ_rows(occ[:n]) == eb.occ_rows(o_ref)
                if chunk == 97 and cap == 4 and plen >= 31 and lf == eb.LF_RANK:
                    assert idx.last_stats()["overflow_patterns"] > 0
            # without the bitmap the flagged patterns are searched as whatever their codes say: only the others must agree
            idx.set_chunk(0)
            counts = np.zeros(len(pats), np.uint32)
            idx.search_into_packed(words, None, plen, len(pats), counts, None, locate=False, lf_variant=lf)
            keep = np.ones(len(pats), bool); keep[[3, 77, 500, 899]] = False
            assert counts[keep].tolist() == c_ref[keep].tolist()


def test_occurrence_overflow_halves_the_chunk(eb, tmp_path, monkeypatch):
    """More occurrences in one chunk than the 32-bit result offsets can address (here: than EDSBWT_OCC_LIMIT): the
    host-buffer calls and the csv call halve the chunk and go on; the device-pointer call reports EDSBWT_E_UNSUPPORTED."""
    rng = np.random.default_rng(77)
    segs = E.random_eds(rng, n_seg=800, max_var=4, max_len=6, p_empty=0.15, nondeg_len=(1, 60))
    _, base = _index_files(tmp_path, segs)
    pats = E.sample_patterns(rng, segs, 400, [2, 3, 4, 8, 16])
    data, off = eb.pack_patterns(pats)
    with eb.Index(base) as idx:
        c_ref, o_ref = idx.search(pats, locate=True)
        csv_ref = str(tmp_path / "ref.csv")
        idx.search_to_csv(pats, csv_ref)
    biggest = int(c_ref.max())
    assert len(o_ref) > 20 * biggest
    # packed batches are halved at multiples of 32 patterns: the limit must admit the heaviest block of 32 length-3 patterns
    w3 = [int(c_ref[i]) for i, p in enumerate(pats) if len(p) == 3]
    limit = max(3 * biggest, max(sum(w3[i:i + 32]) for i in range(0, len(w3), 32)))
    assert len(o_ref) > 3 * limit
    monkeypatch.setenv("EDSBWT_OCC_LIMIT", str(limit))
    with eb.Index(base) as idx:
        c, o = idx.search(pats, locate=True)
        assert c.tolist() == c_ref.tolist() and eb.occ_rows(o) == eb.occ_rows(o_ref)
        assert np.all(np.diff(o["pat"].astype(np.int64)) >= 0)
        fixed = [p for p in pats if len(p) == 3]
        fd, _ = eb.pack_patterns(fixed)
        cf = np.zeros(len(fixed), np.uint32); of = np.zeros(len(o_ref), eb.OCC_DTYPE)
        nf = idx.search_into_fixed(fd, 3, cf, of, locate=True)
        want = [c_ref[i] for i, p in enumerate(pats) if len(p) == 3]
        assert cf.tolist() == want and nf == sum(want)
        words, invalid, _ = idx.pack_patterns(fd, 3)
        cp = np.zeros(len(fixed), np.uint32); op = np.zeros(len(o_ref), eb.OCC_DTYPE)
        assert idx.search_into_packed(words, invalid, 3, len(fixed), cp, op, locate=True) == nf and cp.tolist() == want
        assert eb.occ_rows(op[:nf]) == eb.occ_rows(of[:nf])
        csv = str(tmp_path / "halved.csv")
        idx.search_to_csv(pats, csv)
        assert open(csv, "rb").read() == open(csv_ref, "rb").read()
        import torch
        d_data = torch.from_numpy(data).cuda(); d_off = torch.from_numpy(off.view(np.int64)).cuda()
        with pytest.raises(eb.EdsbwtError) as ei:
            idx.search_device(d_data.data_ptr(), d_off.data_ptr(), len(pats), int(data.size), 16, locate=True)
        assert ei.value.code == -6
        r = idx.search_device(d_data.data_ptr(), d_off.data_ptr(), len(pats), int(data.size), 16, locate=False)     # counting needs no offsets
        assert r.fetch()[0].tolist() == c_ref.tolist()
        r.free()
        idx.set_engine(eb.ENGINE_ITEMS)                              # the packing pass sizes its buffer from total_bytes:
        with pytest.raises(eb.EdsbwtError) as ei:                    # it must cover the offsets' span
            idx.search_device(d_data.data_ptr(), d_off.data_ptr(), len(pats), 64, 0, locate=False)     # (longest pattern unknown: the batch is packed by the separate pass, which sizes its buffer from total_bytes)
        assert ei.value.code == -5


# ---- count-only trace (RECOVERBW=0 stderr line) and the two drop-in executables ---------------------------------
@pytest.mark.parametrize("name", golden_names())
@pytest.mark.parametrize("seed_k,engine,cap", [(-1, 0, 0), (1, 0, 0), (4, 0, 4), (3, 1, 0), (-1, 2, 0), (4, 2, 4), (-1, 3, 0), (1, 3, 0), (4, 3, 4)])
def test_trace_interrupted_positions(eb, tmp_path, name, seed_k, engine, cap):
    g = load_golden(name)
    _, base = _index_files(tmp_path, E.parse_eds(g["eds"]))
    with eb.Index(base, seed_k=seed_k, engine=engine, fused_cap=cap) as idx:
        counts, pos = idx.search_trace(g["patterns"])
    assert counts.tolist() == g["counts"]
    assert pos.tolist() == g["interrupted"]


def test_trace_vs_port_random(eb, oracle_built, tmp_path):
    rng = np.random.default_rng(31)
    segs = E.random_eds(rng, n_seg=2500, max_var=5, max_len=6, p_empty=0.2, nondeg_len=(1, 50))
    _, base = _index_files(tmp_path, segs)
    pats = E.sample_patterns(rng, segs, 800, [1, 2, 3, 5, 8, 13, 21, 32, 50, 64])
    pf = str(tmp_path / "p.txt")
    E.write_patterns(pf, pats)
    c_port, _, st = oracle_built.run_port(base, pf, locate=False)
    for seed_k, cap, engine in [(-1, 0, 0), (5, 4, 0), (2, 0, 0), (5, 4, 2), (-1, 0, 3), (5, 4, 3), (2, 0, 3)]:
        with eb.Index(base, seed_k=seed_k, fused_cap=cap, engine=engine) as idx:
            counts, pos = idx.search_trace(pats)
        assert counts.tolist() == c_port
        assert pos.tolist() == st["interrupted"]


def _run_cli(exe, base, pf, env=None):
    import subprocess
    e = dict(os.environ)
    e.update(env or {})
    return subprocess.run([exe, base, pf], capture_output=True, text=True, env=e)


@pytest.mark.parametrize("name", golden_names())
def test_cli_edsbwtsearch_dropin(eb, tmp_path, name):
    """EDSBWTsearch <index> <patterns>: exit status 1, "Found N" lines, <patterns>output.csv rows, count_found lines;
    EDSBWT_RECOVERBW=0: no csv, "interrupted in position" lines (the reference's -DRECOVERBW=0 build)."""
    import re
    g = load_golden(name)
    _, base = _index_files(tmp_path, E.parse_eds(g["eds"]))
    pf = str(tmp_path / "kmers.txt")
    E.write_patterns(pf, g["patterns"])
    exe = os.path.join(os.path.dirname(eb.__file__), "EDSBWTsearch")
    r = _run_cli(exe, base, pf)
    assert r.returncode == 1, r.stderr
    assert [int(x) for x in re.findall(r"Found (\d+) starting positions of occurrences of pattern number \d+", r.stdout)] == g["counts"]
    assert re.search(r"bs took:[0-9.e+-]+$", r.stdout)
    nfound = sum(1 for c in g["counts"] if c > 0)
    assert "count_found = %d\n" % nfound in r.stderr and "count_not_found = %d\n" % (len(g["counts"]) - nfound) in r.stderr
    assert "The csv file is ready!" in r.stderr and r.stderr.endswith("The End!\n")
    header, rows = E.read_csv_rows(pf + "output.csv")
    assert header == "#Pat\t$_i\tD[i]\tS_j\tS_j[r]\n"
    assert rows == g["rows_reference_order"]          # the reference's row ORDER too: an unsorted diff passes
    os.remove(pf + "output.csv")
    r0 = _run_cli(exe, base, pf, {"EDSBWT_RECOVERBW": "0"})
    assert r0.returncode == 1 and not os.path.exists(pf + "output.csv")
    assert [int(x) for x in re.findall(r"interrupted in position (\d+)", r0.stderr)] == [v for v in g["interrupted"] if v]
    assert "The search is finished!" in r0.stderr
    # usage error: exit(1) with the usage text on stderr (mainEDS-BWT.cpp:42-48)
    import subprocess
    u = subprocess.run([exe, base], capture_output=True, text=True)
    assert u.returncode == 1 and u.stderr.startswith("usage: ")
