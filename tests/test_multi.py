"""Multi-GPU search inside one process (edsbwt_search_multi, csrc/multi.cu): sharded patterns, NCCL gather of counts and tuples
overlapped with the chunked search -- against the single-GPU call on the same batch.  With one visible GPU the sharding,
chunking and segment logic still run (one shard, local copies); the NCCL path needs two (gpurun --gpus 2)."""
import importlib
import os

import numpy as np
import pytest

from oracle import eds_index as E

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def eb():
    return importlib.import_module("eds-bwt_b200")


def _ngpu():
    import torch
    return torch.cuda.device_count()


def _case(tmp_path, seed=123, npat=3000):
    rng = np.random.default_rng(seed)
    segs = E.random_eds(rng, n_seg=2500, max_var=5, max_len=6, p_empty=0.2, nondeg_len=(1, 50))
    ix = E.build_index(segs)
    base = str(tmp_path / "idx")
    E.write_index(ix, base)
    pats = E.sample_patterns(rng, segs, npat, [3, 8, 16, 32, 50])
    return base, pats


@pytest.mark.parametrize("ndev", [1, 2, 4])
def test_multi_equals_single(eb, tmp_path, ndev, monkeypatch):
    if _ngpu() < ndev:
        pytest.skip("needs %d GPUs" % ndev)
    base, pats = _case(tmp_path)
    with eb.Index(base) as idx:
        c_ref, o_ref = idx.search(pats, locate=True)
        csv_ref = str(tmp_path / "single.csv")
        idx.search_to_csv(pats, csv_ref)
    for p2p in ([None] if ndev == 1 else [None, "1"]):
        if p2p:
            monkeypatch.setenv("EDSBWT_MULTI_P2P", p2p)
        else:
            monkeypatch.delenv("EDSBWT_MULTI_P2P", raising=False)
        with eb.MultiIndex(base, devices=list(range(ndev))) as mi:
            assert mi.devices == list(range(ndev)) and (mi.nccl == (ndev > 1 and not p2p))
            for chunks in [1, 3, 7]:
                c, o, info = mi.search(pats, locate=True, chunks_per_gpu=chunks)
                assert c.tolist() == c_ref.tolist() and info["segments"] == ndev * chunks
                assert eb.occ_rows(o) == eb.occ_rows(o_ref)
                assert np.all(np.diff(o["pat"].astype(np.int64)) >= 0)      # ascending pattern order across shards and chunks
                c0, o0, _ = mi.search(pats, locate=False, chunks_per_gpu=chunks)
                assert c0.tolist() == c_ref.tolist() and len(o0) == 0
            # reference row order + csv written from the gathered result == the single-GPU csv, byte for byte
            mi.set_row_order(True)
            csv = str(tmp_path / ("multi_%d_%s.csv" % (ndev, p2p)))
            mi.search(pats, locate=True, chunks_per_gpu=2, csv_path=csv)
            assert open(csv, "rb").read() == open(csv_ref, "rb").read()
            # equal-length batch without an offset array
            fixed = [p for p in pats if len(p) == 16]
            data = np.frombuffer(b"".join(fixed), np.uint8)
            cf, of, _ = mi.search(data, locate=True, fixed_len=16, chunks_per_gpu=2)
            want = [c_ref[i] for i, p in enumerate(pats) if len(p) == 16]
            assert cf.tolist() == want and len(of) == sum(want)


def test_multi_cli(eb, tmp_path):
    """EDSBWTsearch with EDSBWT_DEVICES=all: same csv and report as the single-GPU run"""
    import subprocess
    base, pats = _case(tmp_path, seed=5, npat=500)
    pf = str(tmp_path / "p.txt")
    E.write_patterns(pf, pats)
    exe = os.path.join(os.path.dirname(eb.__file__), "EDSBWTsearch")
    r1 = subprocess.run([exe, base, pf], capture_output=True, text=True)
    csv1 = open(pf + "output.csv", "rb").read()
    os.remove(pf + "output.csv")
    env = dict(os.environ); env["EDSBWT_DEVICES"] = "all"; env["EDSBWT_CHUNKS"] = "3"
    r2 = subprocess.run([exe, base, pf], capture_output=True, text=True, env=env)
    assert r1.returncode == 1 and r2.returncode == 1, r2.stderr
    assert open(pf + "output.csv", "rb").read() == csv1
    f1 = [l for l in r1.stdout.splitlines() if l.startswith("Found ")]
    f2 = [l for l in r2.stdout.splitlines() if l.startswith("Found ")]
    assert f1 == f2 and len(f1) == len(pats)
