"""Threading contract of include/edsbwt.h: calls on ONE handle from several host threads are serialised inside
the library (ctypes drops the GIL during the call, so the threads really contend for the handle's chunk pipeline)
and every thread gets exactly the result of the same call made alone.  The reference object is not re-entrant at
all (per-pattern state in members, EDSBWTsearch.hpp:61-63); a handle must at least be safe."""
import importlib
import threading

import numpy as np
import pytest

from oracle import eds_index as E

pytestmark = pytest.mark.gpu


def test_concurrent_calls_on_one_handle(tmp_path):
    eb = importlib.import_module("eds-bwt_b200")
    rng = np.random.default_rng(71)
    segs = E.random_eds(rng, n_seg=1500, max_var=4, max_len=6, p_empty=0.2, nondeg_len=(1, 40))
    base = str(tmp_path / "idx")
    E.write_index(E.build_index(segs), base)
    shards = [E.sample_patterns(rng, segs, 700, [3, 8, 16, 32, 64]) for _ in range(4)]
    with eb.Index(base, seed_k=4) as idx:
        idx.set_chunk(97)                                     # many chunks per call: the pipeline slots are reused constantly
        alone = [idx.search(p, locate=True) for p in shards]
        got = [None] * len(shards)
        errs = []

        def work(i):
            try:
                for _ in range(3):
                    got[i] = idx.search(shards[i], locate=True)
            except Exception as e:                            # surfaced below: an exception in a thread must fail the test
                errs.append(e)

        ts = [threading.Thread(target=work, args=(i,)) for i in range(len(shards))]
        for t in ts:
            t.start()
        for t in ts:
            t.join()
        assert not errs, errs
        for (c0, o0), (c1, o1) in zip(alone, got):
            assert np.array_equal(c0, c1)
            assert eb.occ_rows(o0) == eb.occ_rows(o1)
