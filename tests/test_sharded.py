"""Multi-GPU plumbing: pattern sharding + result gather.  The collectives are exercised on CPU tensors over
gloo with world sizes 2 and 3 (the local search is played by the oracle -- tests only); on a GPU box the same code runs
over NCCL through ShardedSearcher (world size = number of visible GPUs, at least 1)."""
import importlib
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle import eds_index as E

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def test_shard_range_partitions():
    sh = importlib.import_module("eds-bwt_b200.sharded")
    for n in [0, 1, 2, 7, 8, 9, 1000003]:
        for w in [1, 2, 3, 4, 8]:
            r = [sh.shard_range(n, w, g) for g in range(w)]
            assert r[0][0] == 0 and r[-1][1] == n
            assert all(r[g][1] == r[g + 1][0] for g in range(w - 1))
            assert max(e - b for b, e in r) - min(e - b for b, e in r) <= 1
            assert sh.shard_pad(n, w) == max(e - b for b, e in r)


def _gloo_worker(rank, world, port, base, pf, npat, q):
    sys.path.insert(0, ROOT)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        sh = importlib.import_module("eds-bwt_b200.sharded")
        from oracle import run as R
        b, e = sh.shard_range(npat, world, rank)
        # local "search" of this rank's shard: the oracle on lines [b, e) of the pattern file
        lines = open(pf, "rb").read().split(b"\n")[:npat]
        spf = "%s.%d" % (pf, rank)
        with open(spf, "wb") as f:
            f.write(b"\n".join(lines[b:e]))
        if e > b:
            counts, rows, _ = R.run_port(base, spf, locate=True)
        else:
            counts, rows = [], []
        c_loc = torch.tensor(np.array(counts, dtype=np.uint32).view(np.int32), dtype=torch.int32)
        o = np.array([(p - 1 + b, w, s, j, r) for (p, w, s, j, r) in rows], dtype=np.uint32).reshape(-1, 5)
        o_loc = torch.from_numpy(o.view(np.int32).copy())
        allc, occ, tots = sh.gather_results(c_loc, o_loc, npat, None, dst=0)
        q.put((rank, allc.numpy().view(np.uint32).tolist(), None if occ is None else occ.numpy().view(np.uint32).tolist(), tots))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,npat", [(2, 41), (3, 10), (2, 1)])
def test_gather_over_gloo(world, npat, oracle_built, tmp_path):
    rng = np.random.default_rng(5 + world)
    segs = E.random_eds(rng, n_seg=200, max_var=4, max_len=5, p_empty=0.2)
    ix = E.build_index(segs)
    base = str(tmp_path / "idx")
    E.write_index(ix, base)
    pats = E.sample_patterns(rng, segs, npat, [2, 3, 5, 8, 12])
    pf = str(tmp_path / "p.txt")
    E.write_patterns(pf, pats)
    c_ref, rows_ref, _ = oracle_built.run_port(base, pf, locate=True)
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_gloo_worker, args=(r, world, port, base, pf, npat, q)) for r in range(world)]
    for p in procs:
        p.start()
    got = [q.get(timeout=120) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    got.sort()
    for rank, counts, occ, tots in got:
        assert counts == c_ref                                  # every rank holds the counts of the whole batch
        assert sum(tots) == len(rows_ref)
        if rank == 0:
            rows = [(p + 1, w, s, j, r) for (p, w, s, j, r) in occ]
            assert sorted(rows) == sorted(rows_ref)
            assert [r[0] for r in rows] == sorted(r[0] for r in rows)      # grouped by ascending pattern without sorting
        else:
            assert occ is None


def _gloo_chunk_worker(rank, world, port, base, pf, npat, nchunks, q):
    sys.path.insert(0, ROOT)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        sh = importlib.import_module("eds-bwt_b200.sharded")
        from oracle import run as R
        size_group = dist.new_group(backend="gloo")
        b, e = sh.shard_range(npat, world, rank)
        lines = open(pf, "rb").read().split(b"\n")[:npat]
        g = sh.ChunkedGather(npat, nchunks, None, size_group, dst=0)
        for c in range(nchunks):
            cb, ce = sh.chunk_range(b, e, nchunks, c)
            counts, rows = [], []
            if ce > cb:                                     # local "search" of the chunk: the oracle on its lines
                spf = "%s.%d.%d" % (pf, rank, c)
                with open(spf, "wb") as f:
                    f.write(b"\n".join(lines[cb:ce]))
                counts, rows, _ = R.run_port(base, spf, locate=True)
            c_loc = torch.tensor(np.array(counts, dtype=np.uint32).view(np.int32), dtype=torch.int32)
            o = np.array([(p - 1 + cb, w, s, j, r) for (p, w, s, j, r) in rows], dtype=np.uint32).reshape(-1, 5)
            g.post(c, c_loc, torch.from_numpy(o.view(np.int32).copy()))
        allc, segs = g.finish()
        if rank == 0:
            occ = np.concatenate([s_.numpy().view(np.uint32).reshape(-1, 5) for s_ in segs]) if segs else np.zeros((0, 5), np.uint32)
            q.put((rank, allc.numpy().view(np.uint32).tolist(), occ.tolist(), len(segs)))
        else:
            q.put((rank, None, None, 0))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,npat,nchunks", [(2, 41, 3), (3, 10, 4), (2, 5, 1)])
def test_chunked_gather_over_gloo(world, npat, nchunks, oracle_built, tmp_path):
    """the chunk-overlapped gather of bench.py's multi-GPU leg: counts of the whole batch + tuple segments in ascending
    pattern order on rank 0 (chunks smaller than the world, empty chunks included)"""
    rng = np.random.default_rng(50 + world)
    segs = E.random_eds(rng, n_seg=200, max_var=4, max_len=5, p_empty=0.2)
    ix = E.build_index(segs)
    base = str(tmp_path / "idx")
    E.write_index(ix, base)
    pats = E.sample_patterns(rng, segs, npat, [2, 3, 5, 8, 12])
    pf = str(tmp_path / "p.txt")
    E.write_patterns(pf, pats)
    c_ref, rows_ref, _ = oracle_built.run_port(base, pf, locate=True)
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_gloo_chunk_worker, args=(r, world, port, base, pf, npat, nchunks, q)) for r in range(world)]
    for p in procs:
        p.start()
    got = [q.get(timeout=120) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    for rank, counts, occ, nseg in got:
        if rank == 0:
            assert counts == c_ref and nseg == world * nchunks
            rows = [(p + 1, w, s, j, r) for (p, w, s, j, r) in occ]
            assert sorted(rows) == sorted(rows_ref)
            assert [r[0] for r in rows] == sorted(r[0] for r in rows)      # ascending pattern order straight from the segments


# ---- GPU: the real thing over NCCL -------------------------------------------------------------------------------------
def _nccl_worker(rank, world, port, base, pats, q):
    sys.path.insert(0, ROOT)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    try:
        eb = importlib.import_module("eds-bwt_b200")
        sh = importlib.import_module("eds-bwt_b200.sharded")
        data, off = eb.pack_patterns(pats)
        with eb.Index(base, device=rank) as idx:
            s = sh.ShardedSearcher(idx)
            out = []
            for lf in (eb.LF_RANK, eb.LF_MOVE):
                counts, occ = s.search(data, off, locate=True, lf_variant=lf)
                out.append((counts.tolist(), None if occ is None else eb.occ_rows(occ)))
            # the gather over peer memory: every chunk of the shard pushed into rank 0's arena (CUDA IPC, NVLink peer copies)
            npat = len(off) - 1
            b, e = sh.shard_range(npat, world, rank)
            for nchunks in (1, 3):
                g = sh.PeerGather(eb, idx, npat, nchunks, len(out[0][1]) + 64 + rank if out[0][1] is not None else 1, None, 0)   # (capacities differ per rank: the largest wins)
                for rep in range(2):                                    # the arena is reused batch after batch
                    g.begin()
                    for c in range(nchunks):
                        cb, ce = sh.chunk_range(b, e, nchunks, c)
                        d_data = torch.from_numpy(np.ascontiguousarray(data[int(off[cb]):int(off[ce])]) if off[ce] > off[cb] else np.zeros(1, np.uint8)).cuda()
                        d_off = torch.from_numpy((off[cb:ce + 1] - off[cb]).astype(np.int64)).cuda()
                        r = idx.search_device(d_data.data_ptr(), d_off.data_ptr(), ce - cb, int(off[ce] - off[cb]), 0, locate=True, pat_base=cb)
                        g.push(c, r)
                    allc, segs = g.finish()
                if rank == 0:
                    occ = torch.cat(segs).cpu().numpy().view(np.uint32).reshape(-1).view(eb.OCC_DTYPE) if segs else np.zeros(0, eb.OCC_DTYPE)
                    assert len(segs) == world * nchunks and np.all(np.diff(occ["pat"].astype(np.int64)) >= 0)
                    out.append((allc.cpu().numpy().view(np.uint32).tolist(), eb.occ_rows(occ)))
                else:
                    out.append((out[0][0], None))
                dist.barrier()
                g.close()
        q.put((rank, out))
    finally:
        dist.destroy_process_group()


@pytest.mark.gpu
def test_sharded_search_nccl(oracle_built, tmp_path):
    world = min(torch.cuda.device_count(), 2)
    assert world >= 1
    rng = np.random.default_rng(17)
    segs = E.random_eds(rng, n_seg=1500, max_var=4, max_len=6, p_empty=0.15, nondeg_len=(1, 40))
    ix = E.build_index(segs)
    base = str(tmp_path / "idx")
    E.write_index(ix, base)
    pats = E.sample_patterns(rng, segs, 501, [4, 9, 16, 32, 33])
    pf = str(tmp_path / "p.txt")
    E.write_patterns(pf, pats)
    c_ref, rows_ref, _ = oracle_built.run_port(base, pf, locate=True)
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_nccl_worker, args=(r, world, port, base, pats, q)) for r in range(world)]
    for p in procs:
        p.start()
    got = [q.get(timeout=300) for _ in range(world)]
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    for rank, out in got:
        for counts, rows in out:
            assert counts == c_ref
            if rank == 0:
                assert rows == sorted(rows_ref)
            else:
                assert rows is None
