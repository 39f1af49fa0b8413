"""CPU tests of the drop-in boundary: the C-ABI library loads, exports every symbol include/edsbwt.h declares,
its host-only entry points work, and compute entry points fail loudly without a GPU (no CPU fallback)."""
import ctypes as C
import importlib
import os
import re

import numpy as np
import pytest

from oracle import eds_index as E, sdsl_io as S

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))


@pytest.fixture(scope="module")
def eb():
    import sys
    sys.path.insert(0, ROOT)
    import __graft_entry__ as G
    G.build()
    return importlib.import_module("eds-bwt_b200")


def test_header_symbols_exported(eb):
    hdr = open(os.path.join(ROOT, "include", "edsbwt.h")).read()
    declared = set(re.findall(r"\b(edsbwt_[a-z_0-9]+)\s*\(", hdr))
    assert declared == set(eb.EXPORTS)
    L = eb.lib()
    for name in declared:
        assert hasattr(L, name), name


def test_pattern_reader_and_csv(eb, tmp_path):
    pf = str(tmp_path / "p.txt")
    with open(pf, "wb") as f:
        f.write(b"TATT\nACT\n\nTTAT")
    data, off = eb.read_patterns(pf)
    assert off.tolist() == [0, 4, 7, 7, 11] and bytes(data) == b"TATTACTTTAT"
    occ = np.array([(0, 3, 2, 0, 0), (2, 0, 1, 0, 1)], dtype=eb.OCC_DTYPE)
    out = str(tmp_path / "o.csv")
    eb.write_csv(out, occ)
    assert open(out).read() == "#Pat\t$_i\tD[i]\tS_j\tS_j[r]\n1\t3\t2\t0\t0\n3\t0\t1\t0\t1\n"
    eb.write_csv(out, occ[:1], pat_base=10, move=True)
    assert open(out).read() == "#Pat\t$_i\tD[i]\tS_j\tS_j[r] \n11\t3\t2\t0\t0\n"


def _pack_reference(pats, plen, letters=b"ACGT"):
    """numpy-free restatement of the packed layout of include/edsbwt.h"""
    stride = (plen + 31) // 32
    words, bad = [], []
    for i, p in enumerate(pats):
        ok = all(ch in letters for ch in p)
        for w in range(stride):
            hi = plen - 32 * w
            lo = max(0, hi - 32)
            v = 0
            for j in range(lo, hi):
                c = letters.index(p[j]) if p[j] in letters else 0
                v |= (c & 3) << (2 * (j - lo))
            words.append((v << (2 * (32 - (hi - lo)))) & (2 ** 64 - 1) if ok else None)
        if not ok:
            bad.append(i)
    return words, bad


@pytest.mark.parametrize("plen", [1, 5, 31, 32, 33, 64, 100])
@pytest.mark.parametrize("nthreads", [1, 4])
def test_host_packer(eb, plen, nthreads):
    """edsbwt_pack_patterns_alphabet (host-only): words and the invalid bitmap against the layout's definition"""
    rng = np.random.default_rng(plen)
    npat = 20000 if nthreads > 1 else 3000
    data = rng.choice(np.frombuffer(b"ACGT", np.uint8), size=npat * plen)
    for i in rng.integers(0, npat, 7):
        data[i * plen + rng.integers(0, plen)] = ord("N")
    pats = [bytes(data[i * plen:(i + 1) * plen]) for i in range(npat)]
    words, invalid, nbad = eb.pack_patterns_alphabet(b"#ACGT", data, plen, nthreads)
    exp, bad = _pack_reference(pats, plen)
    assert nbad == len(bad)
    assert [i for i in range(npat) if (invalid[i >> 5] >> (i & 31)) & 1] == bad
    got = words.tolist()
    assert all(e is None or e == g for e, g in zip(exp, got))
    # a three-letter alphabet: T is not a letter
    w3, inv3, nbad3 = eb.pack_patterns_alphabet(b"#ACG", data[:64 * plen], plen, 1)
    assert nbad3 == sum(1 for p in pats[:64] if any(ch not in b"ACG" for ch in p))


def test_no_cpu_fallback(eb, tmp_path):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    ix = E.build_index(E.parse_eds(b"{TT,TA,C}{T}{TT,A,G}{TT,}{ATT}"))
    base = str(tmp_path / "t")
    E.write_index(ix, base)
    with pytest.raises(eb.EdsbwtError) as ei:
        eb.Index(base)
    assert ei.value.code == -4
    with pytest.raises(eb.EdsbwtError):
        eb.gather_bench(0, 1 << 20, 1 << 10, 1)


def test_host_entry_points_report_io_errors(eb, tmp_path):
    """Nothing may cross the C ABI as an exception: unreadable / unwritable paths come back as EDSBWT_E_IO (-1)
    with a message, the reference's `exit(1)` paths (EDSBWTsearch.cpp:183-185, 200-203) turned into return codes."""
    L = eb.lib()
    d = str(tmp_path)                                   # a directory is not a pattern file
    b, o, n = C.c_void_p(), C.c_void_p(), C.c_uint32(7)
    for bad in (d, os.path.join(d, "missing.txt")):
        rc = L.edsbwt_read_patterns(bad.encode(), C.byref(b), C.byref(o), C.byref(n))
        assert rc == -1 and n.value == 7 and not b and not o
        assert os.path.basename(bad) in L.edsbwt_last_error().decode()
    src = str(tmp_path / "in.eds")
    open(src, "wb").write(b"{AC}{G,T}{A}")
    assert L.edsbwt_string_check(src.encode(), os.path.join(d, "no_such_dir", "out.eds").encode()) == -1
    assert L.edsbwt_string_check(d.encode(), os.path.join(d, "out.eds").encode()) == -1
    occ = np.zeros(1, dtype=eb.OCC_DTYPE)
    rc = L.edsbwt_write_csv(os.path.join(d, "no_such_dir", "o.csv").encode(), occ.ctypes.data, 1, 0, 0, 0)
    assert rc == -1 and "o.csv" in L.edsbwt_last_error().decode()
    if os.path.exists("/dev/full"):                     # a write that fails after a successful open
        rc = L.edsbwt_write_csv(b"/dev/full", occ.ctypes.data, 1, 0, 0, 0)
        assert rc == -1


def test_empty_pattern_file(eb, tmp_path):
    pf = str(tmp_path / "empty.txt")
    open(pf, "wb").close()
    data, off = eb.read_patterns(pf)
    assert off.tolist() == [0] and len(data) == 0


def test_struct_layouts_match_the_header(eb, tmp_path):
    """the ctypes mirrors of include/edsbwt.h's structs (tests / bench binding) have the C compiler's sizes and field offsets"""
    import ctypes as C
    import subprocess
    structs = {"edsbwt_info": eb.Info, "edsbwt_open_opts": eb.OpenOpts, "edsbwt_stats": eb.Stats, "edsbwt_gather_cfg": eb.GatherCfg}
    lines = ['#include <stdio.h>', '#include <stddef.h>', '#include "edsbwt.h"', 'int main(void) {']
    for cname, cls in structs.items():
        lines.append('printf("%s sizeof %%zu\\n", sizeof(%s));' % (cname, cname))
        for fname, _ in cls._fields_:
            cfield = {"granule_bytes": "granule_bytes"}.get(fname, fname)
            lines.append('printf("%s %s %%zu\\n", offsetof(%s, %s));' % (cname, fname, cname, cfield))
    lines.append('printf("edsbwt_occ sizeof %zu\\n", sizeof(edsbwt_occ)); return 0; }')
    src = tmp_path / "layout.c"
    src.write_text("\n".join(lines))
    exe = str(tmp_path / "layout")
    inc = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "include")
    subprocess.run(["gcc", "-std=c99", "-I", inc, str(src), "-o", exe], check=True)
    out = subprocess.run([exe], check=True, capture_output=True, text=True).stdout.split("\n")
    seen = 0
    for ln in out:
        t = ln.split()
        if len(t) != 3:
            continue
        cname, what, val = t[0], t[1], int(t[2])
        if cname == "edsbwt_occ":
            assert val == eb.OCC_DTYPE.itemsize == 20
        elif what == "sizeof":
            assert C.sizeof(structs[cname]) == val, cname
        else:
            assert getattr(structs[cname], what).offset == val, (cname, what)
        seen += 1
    assert seen > 40
