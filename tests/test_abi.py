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
