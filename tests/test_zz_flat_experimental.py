"""Parity of the EXPERIMENTAL warp-flat first pass of the fused kernel (EDSBWT_ENGINE_FUSED_FLAT)
against the oracle.  The kernel was written at the end of round 1 without a GPU at hand: it is opt-in in the library and
these tests are opt-in too (EDSBWT_TEST_FLAT=1), so that the default `pytest -m gpu` run judges only measured code."""
import importlib
import os

import numpy as np
import pytest

from conftest import golden_names, load_golden
from oracle import eds_index as E

pytestmark = [pytest.mark.gpu,
              pytest.mark.skipif(os.environ.get("EDSBWT_TEST_FLAT") != "1", reason="experimental kernel: set EDSBWT_TEST_FLAT=1")]


@pytest.fixture(scope="module", params=["1", "2", "3"])
def eb(request):
    """EDSBWT_FUSED_FLAT variants (fused.cu): 1 = warp-flat fetch at 10 CTAs/SM (the one run green at the end of round 1),
    2 = the same compiled for 12 CTAs/SM, 3 = + warp-wide link gather.  The variable is read when an index is opened."""
    old = os.environ.get("EDSBWT_FUSED_FLAT")
    os.environ["EDSBWT_FUSED_FLAT"] = request.param
    yield importlib.import_module("eds-bwt_b200")
    if old is None:
        os.environ.pop("EDSBWT_FUSED_FLAT", None)
    else:
        os.environ["EDSBWT_FUSED_FLAT"] = old


@pytest.mark.parametrize("name", golden_names())
@pytest.mark.parametrize("seed_k,cap", [(-1, 0), (1, 0), (2, 5), (3, 14)])
def test_flat_golden(eb, tmp_path, name, seed_k, cap):
    g = load_golden(name)
    base = str(tmp_path / "idx")
    E.write_index(E.build_index(E.parse_eds(g["eds"])), base)
    with eb.Index(base, seed_k=seed_k, engine=eb.ENGINE_FUSED, fused_cap=cap) as idx:        # flat via the environment variable
        counts, occ = idx.search(g["patterns"], locate=True)
        assert counts.tolist() == g["counts"]
        assert eb.occ_rows(occ) == sorted(g["rows_reference_order"])
        _, pos = idx.search_trace(g["patterns"])
        assert pos.tolist() == g["interrupted"]


@pytest.mark.parametrize("seed,kw,seed_k,cap", [
    (31, dict(n_seg=300, max_var=5, max_len=5, p_empty=0.3), 2, 0),
    (32, dict(n_seg=2000, max_var=4, max_len=6, p_empty=0.15, nondeg_len=(1, 40)), 4, 0),
    (33, dict(n_seg=1200, max_var=8, max_len=3, p_empty=0.45, nondeg_len=(1, 3)), 3, 0),     # '#' everywhere: queue overflows
    (34, dict(n_seg=5000, max_var=3, max_len=8, p_empty=0.1, nondeg_len=(5, 120)), -1, 6),
])
def test_flat_random_vs_port(eb, oracle_built, tmp_path, seed, kw, seed_k, cap):
    rng = np.random.default_rng(seed)
    segs = E.random_eds(rng, **kw)
    base = str(tmp_path / "idx")
    E.write_index(E.build_index(segs), base)
    pats = E.sample_patterns(rng, segs, 3000, [1, 2, 3, 5, 8, 13, 21, 32, 50, 64, 100]) + [b"ACGNT", b"#A", b"A" * 300]
    pf = str(tmp_path / "p.txt")
    E.write_patterns(pf, pats)
    with eb.Index(base, seed_k=seed_k, engine=eb.ENGINE_FUSED, fused_cap=cap) as idx:        # flat via the environment variable
        k = idx.info().seed_k
        c_port, rows_port, st_port = oracle_built.run_port(base, pf, locate=True, k=k)
        counts, occ = idx.search(pats, locate=True)
        assert counts.tolist() == c_port
        assert eb.occ_rows(occ) == sorted(rows_port)
        idx.set_stats(True, canonical_k=k)                     # counting instantiation of the flat loop: byte-model counters
        counts1, occ1 = idx.search(pats, locate=True)
        st = idx.last_stats()
        assert counts1.tolist() == c_port and eb.occ_rows(occ1) == sorted(rows_port)
        for key in "ISELO":
            assert st[key] == st_port[key], (key, st[key], st_port[key])
    variant = os.environ.pop("EDSBWT_FUSED_FLAT")
    try:
        with eb.Index(base, seed_k=seed_k, engine=eb.ENGINE_FUSED, fused_cap=cap) as idx:      # and the shipped kernel agrees
            counts0, occ0 = idx.search(pats, locate=True)
            assert np.array_equal(counts0, counts) and eb.occ_rows(occ0) == eb.occ_rows(occ)
    finally:
        os.environ["EDSBWT_FUSED_FLAT"] = variant


def test_flat_engine_option(tmp_path):
    """The open option selects variant 1 without the environment variable."""
    eb = importlib.import_module("eds-bwt_b200")
    variant = os.environ.pop("EDSBWT_FUSED_FLAT", None)
    try:
        g = load_golden("readme_test")
        base = str(tmp_path / "idx")
        E.write_index(E.build_index(E.parse_eds(g["eds"])), base)
        with eb.Index(base, seed_k=1, engine=eb.ENGINE_FUSED_FLAT) as idx:
            counts, occ = idx.search(g["patterns"], locate=True)
            assert counts.tolist() == g["counts"] and eb.occ_rows(occ) == sorted(g["rows_reference_order"])
    finally:
        if variant is not None:
            os.environ["EDSBWT_FUSED_FLAT"] = variant
