"""edsbwt_count_static (csrc/static_scratch.cu): the count-only search that allocates nothing -- its scratch is a file-scope
device array -- against the existing entry point edsbwt_search_device on the same device-resident batches.  Integer work:
the bar is equality.  Collected last on purpose."""
import importlib

import numpy as np
import pytest

from oracle import eds_index as E

pytestmark = pytest.mark.gpu

MAX_PATTERNS = 65536      # EDSBWT_STATIC_MAX_PATTERNS (include/edsbwt.h)


@pytest.fixture(scope="module")
def eb():
    return importlib.import_module("eds-bwt_b200")


def _device_batch(eb, pats):
    import torch
    data, off = eb.pack_patterns(pats)
    d_data = torch.from_numpy(data).cuda()
    d_off = torch.from_numpy(off.view(np.int64)).cuda()
    return data, d_data, d_off


def _both(eb, idx, pats):
    """(counts of the existing entry point, counts of edsbwt_count_static, patterns left unresolved)"""
    import torch
    data, d_data, d_off = _device_batch(eb, pats)
    r = idx.search_device(d_data.data_ptr(), d_off.data_ptr(), len(pats), int(data.size), 0, locate=False)
    want = r.fetch()[0]
    r.free()
    d_counts = torch.full((len(pats),), 0x7fffffff, dtype=torch.int32, device="cuda")      # poisoned: every count must be written
    left = idx.count_static(d_data.data_ptr(), d_off.data_ptr(), len(pats), d_counts.data_ptr())
    torch.cuda.synchronize()
    return want, d_counts.cpu().numpy().view(np.uint32), left


@pytest.mark.parametrize("cap", [0, 4])        # 4 slots: part of the batch overflows the first pass and takes the second one
def test_count_static_equals_search_device_two_calls_in_a_row(eb, tmp_path, cap):
    rng = np.random.default_rng(91)
    segs = E.random_eds(rng, n_seg=1500, max_var=5, max_len=5, p_empty=0.25, nondeg_len=(1, 25))
    base = str(tmp_path / "idx")
    E.write_index(E.build_index(segs), base)
    big = E.sample_patterns(rng, segs, 5000, [1, 2, 3, 5, 8, 13, 21, 32, 50, 64, 100]) + [b"ACGNT", b"#A", b"A" * 300]
    small = E.sample_patterns(rng, segs, 700, [4, 9, 16, 33])
    with eb.Index(base, fused_cap=cap) as idx:
        # two calls in a row on the shared scratch (the second batch is smaller: stale contents of the first must not leak),
        # then the first batch again
        for pats in (big, small, big):
            want, got, left = _both(eb, idx, pats)
            assert left == 0
            assert got.tolist() == want.tolist()
        assert sum(want.tolist()) > 0
        # a second handle on the same device shares the array
        with eb.Index(base, seed_k=3) as idx2:
            want, got, left = _both(eb, idx2, small)
            assert left == 0 and got.tolist() == want.tolist()


def test_count_static_rejects_batches_beyond_the_maximum(eb, tmp_path):
    import torch
    rng = np.random.default_rng(92)
    segs = E.random_eds(rng, n_seg=60)
    base = str(tmp_path / "idx")
    E.write_index(E.build_index(segs), base)
    with eb.Index(base) as idx:
        pats = [b"ACGT"] * (MAX_PATTERNS + 1)
        data, d_data, d_off = _device_batch(eb, pats)
        d_counts = torch.zeros(len(pats), dtype=torch.int32, device="cuda")
        with pytest.raises(eb.EdsbwtError) as ei:
            idx.count_static(d_data.data_ptr(), d_off.data_ptr(), len(pats), d_counts.data_ptr())
        assert ei.value.code == -7 and "EDSBWT_STATIC_MAX_PATTERNS" in str(ei.value)
        # exactly the maximum is accepted
        want, got, left = _both(eb, idx, pats[:MAX_PATTERNS])
        assert left == 0 and got.tolist() == want.tolist()
        assert idx.count_static(0, 0, 0, 0) == 0          # an empty batch is a no-op
