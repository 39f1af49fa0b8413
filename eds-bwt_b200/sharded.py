"""Multi-GPU search: patterns sharded across the ranks of one box, index replicated on every GPU.

One process per GPU (torchrun); `torch.distributed` is plumbing only.  The search itself has no exchange step: rank g
searches the contiguous block [g*N/G, (g+1)*N/G) of the pattern list on its own copy of the index.  The only
collectives are the result gather the north star names:

    counts   all_gather_into_tensor of the per-pattern counts (u32, shards padded to ceil(N/G))   -> ncclAllGather
    tuples   per-rank occurrence totals all-gathered (8 B each), then a gather-v of the 20-byte tuples to `dst`
             as one batch of isend/irecv                                                          -> ncclGroupStart{Send/Recv}

Tuples carry batch-global pattern indices (the library adds the shard base), so the gathered buffer is already grouped by
ascending pattern: shards are contiguous and each shard's tuples are grouped by pattern.

The functions below only move tensors; they run unchanged on CPU tensors over gloo (tests/test_sharded.py, world size 2).
"""
import numpy as np
import torch
import torch.distributed as dist


def shard_range(npat, world, rank):
    """Contiguous block of the pattern list owned by `rank`: [rank*N/G, (rank+1)*N/G)."""
    return (npat * rank) // world, (npat * (rank + 1)) // world


def shard_pad(npat, world):
    """Largest shard size = the per-rank slot of the counts all-gather."""
    return max(shard_range(npat, world, r)[1] - shard_range(npat, world, r)[0] for r in range(world)) if world else 0


def gather_results(counts_local, occ_local, npat, group=None, dst=0):
    """counts_local: int32 tensor [shard size] (u32 bit patterns); occ_local: int32 tensor [k, 5] (edsbwt_occ rows).
    -> (counts of the whole batch, int32 [npat], on every rank;
        tuples of the whole batch, int32 [total, 5], on rank `dst` and None elsewhere;
        per-rank occurrence totals, python list)"""
    world = dist.get_world_size(group)
    rank = dist.get_rank(group)
    dev = counts_local.device
    b, e = shard_range(npat, world, rank)
    assert counts_local.numel() == e - b, (counts_local.numel(), b, e)
    pad = shard_pad(npat, world)
    # ---- counts: one all-gather of equal slots
    slot = torch.zeros(pad, dtype=torch.int32, device=dev)
    slot[: e - b] = counts_local
    allc = torch.empty(pad * world, dtype=torch.int32, device=dev)
    dist.all_gather_into_tensor(allc, slot, group=group)
    parts = []
    for r in range(world):
        rb, re_ = shard_range(npat, world, r)
        parts.append(allc[r * pad: r * pad + (re_ - rb)])
    counts = torch.cat(parts) if parts else allc[:0]
    # ---- occurrence totals of every rank (exclusive scan = where each rank's tuples land)
    k = int(occ_local.shape[0])
    tot = torch.tensor([k], dtype=torch.int64, device=dev)
    tots = torch.empty(world, dtype=torch.int64, device=dev)
    dist.all_gather_into_tensor(tots, tot, group=group)
    tots = [int(x) for x in tots.tolist()]
    # ---- gather-v of the tuples to dst
    occ = None
    ops = []
    if rank == dst:
        occ = torch.empty((sum(tots), 5), dtype=torch.int32, device=dev)
        ofs = 0
        for r in range(world):
            if r == dst:
                if tots[r]:
                    occ[ofs: ofs + tots[r]] = occ_local
            elif tots[r]:
                ops.append(dist.P2POp(dist.irecv, occ[ofs: ofs + tots[r]], _global_rank(r, group), group))
            ofs += tots[r]
    elif k:
        ops.append(dist.P2POp(dist.isend, occ_local.contiguous(), _global_rank(dst, group), group))
    if ops:
        for w in dist.batch_isend_irecv(ops):
            w.wait()
    return counts, occ, tots


def chunk_range(b, e, nchunks, c):
    """Chunk c of the shard [b, e): [b + (e-b)*c/C, b + (e-b)*(c+1)/C)."""
    n = e - b
    return b + (n * c) // nchunks, b + (n * (c + 1)) // nchunks


class ChunkedGather:
    """Gather to rank `dst` of a shard that is searched in chunks: the counts and tuples of chunk c are put on the wire as soon
    as the chunk has been searched and travel while chunk c + 1 is searched, so the collective hides behind the kernels
    (round 1 gathered the whole shard after the search: the 7-to-1 incast doubled the step at N = 8).

        g = ChunkedGather(npat, nchunks, group, size_group, dst)
        for c in range(nchunks): g.post(c, counts_c, occ_c)       # right after chunk c has been searched
        counts, segments = g.finish()                               # rank dst: counts of the whole batch + tuple segments

    Sizes are exchanged chunk by chunk over `size_group` (a group of its own -- gloo on the host in bench.py -- so that a size
    exchange never queues behind the bulk transfers on the NCCL stream).  The tuples of (rank r, chunk c) land in their own
    buffer; the list of segments in (r, c) order is the batch in ascending pattern order, without a final copy.  Everything is
    issued as non-blocking isend / irecv (ncclSend / ncclRecv in one group per chunk); nothing waits before finish()."""

    def __init__(self, npat, nchunks, group=None, size_group=None, dst=0):
        self.npat, self.nchunks, self.group, self.dst = npat, nchunks, group, dst
        self.size_group = size_group if size_group is not None else group
        self.world, self.rank = dist.get_world_size(group), dist.get_rank(group)
        self.works, self.keep, self.segs, self.counts = [], [], {}, None
        self.bytes_moved = 0

    def post(self, c, counts_c, occ_c):
        dev = counts_c.device
        k = int(occ_c.shape[0])
        # sizes of this chunk on every rank (8 bytes each; host tensors when size_group is a gloo group)
        sdev = torch.device("cpu") if dist.get_backend(self.size_group) == "gloo" else dev
        tots = torch.empty(self.world, dtype=torch.int64, device=sdev)
        dist.all_gather_into_tensor(tots, torch.tensor([k], dtype=torch.int64, device=sdev), group=self.size_group)
        tots = [int(x) for x in tots.tolist()]
        ops = []
        if self.rank == self.dst:
            if self.counts is None:
                self.counts = torch.empty(self.npat, dtype=torch.int32, device=dev)
            for r in range(self.world):
                b, e = shard_range(self.npat, self.world, r)
                cb, ce = chunk_range(b, e, self.nchunks, c)
                if r == self.dst:
                    self.counts[cb:ce] = counts_c
                    self.segs[(r, c)] = occ_c
                    continue
                if ce > cb:
                    ops.append(dist.P2POp(dist.irecv, self.counts[cb:ce], _global_rank(r, self.group), self.group))
                seg = torch.empty((tots[r], 5), dtype=torch.int32, device=dev)
                self.segs[(r, c)] = seg
                if tots[r]:
                    ops.append(dist.P2POp(dist.irecv, seg, _global_rank(r, self.group), self.group))
                self.bytes_moved += 4 * (ce - cb) + 20 * tots[r]
        else:
            if counts_c.numel():
                ops.append(dist.P2POp(dist.isend, counts_c, _global_rank(self.dst, self.group), self.group))
            if k:
                ops.append(dist.P2POp(dist.isend, occ_c, _global_rank(self.dst, self.group), self.group))
        self.keep.append((counts_c, occ_c))
        if ops:
            self.works += dist.batch_isend_irecv(ops)
        return tots

    def finish(self):
        for w in self.works:
            w.wait()
        self.works, self.keep = [], []
        if self.rank != self.dst:
            return None, None
        order = [(r, c) for r in range(self.world) for c in range(self.nchunks)]
        return self.counts, [self.segs[k] for k in order]


class PeerGather:
    """The gather over peer memory (CUDA IPC + NVLink peer copies; include/edsbwt.h edsbwt_result_push): rank `dst` owns an arena
    in its HBM -- counts[npat] followed by one tuple region of `tuple_cap` entries per rank -- which every other rank maps once.
    A rank pushes chunk c of its shard as soon as it has searched it: the counts to their final place, the tuples behind the
    previous chunk's in its own region.  Nothing is exchanged beforehand and rank dst posts nothing, so a chunk costs its rank two
    cudaMemcpyAsync calls and travels under the search of the next chunk; finish() waits for the rank's own pushes and all-gathers
    the chunk sizes (one small collective per batch), after which rank dst holds the batch: counts + tuple segments in (rank,
    chunk) order = ascending pattern order.

        g = PeerGather(eb, index, npat, nchunks, tuple_cap, group, dst)        # once: arena + IPC handshake
        g.begin(); [g.push(c, device_result_c) ...]; counts, segments = g.finish()   # per batch
    """

    def __init__(self, eb, index, npat, nchunks, tuple_cap, group=None, dst=0):
        self.eb, self.index, self.npat, self.nchunks, self.cap, self.group, self.dst = eb, index, npat, nchunks, int(tuple_cap), group, dst
        self.world, self.rank = dist.get_world_size(group), dist.get_rank(group)
        self.dev = index.device
        if self.world > 1:                            # every rank must lay the arena out the same way: the largest capacity asked for
            t = torch.tensor([self.cap], dtype=torch.int64, device=torch.device("cuda", self.dev))
            dist.all_reduce(t, op=dist.ReduceOp.MAX, group=group)
            self.cap = int(t.item())
        self.tup_off = (4 * npat + 255) & ~255
        nbytes = self.tup_off + 20 * self.cap * self.world
        box = [None]
        if self.rank == dst:
            self.base = eb.device_alloc(self.dev, nbytes)
            box[0] = eb.ipc_export(self.dev, self.base)
        if self.world > 1:
            dist.broadcast_object_list(box, src=_global_rank(dst, group), group=group)
        if self.rank != dst:
            self.base = eb.ipc_open(self.dev, box[0])
        self.b, self.e = shard_range(npat, self.world, self.rank)
        self.live, self.sizes, self.used = [], [], 0
        self.bytes_moved = 0

    def begin(self):
        self.live, self.sizes, self.used = [], [], 0

    def push(self, c, result):
        cb, ce = chunk_range(self.b, self.e, self.nchunks, c)
        assert result.npat == ce - cb and len(self.sizes) == c
        if self.used + result.nocc > self.cap:
            raise RuntimeError("PeerGather: rank %d needs more than its %d tuple slots; raise tuple_cap" % (self.rank, self.cap))
        self.index.push(result, self.base + 4 * cb, self.base + self.tup_off + 20 * (self.rank * self.cap + self.used))
        self.live.append(result)                      # its buffers are the source of the copy until finish()
        self.sizes.append(result.nocc)
        self.used += result.nocc

    def finish(self):
        self.index.push_wait()                        # this rank's chunks have landed
        dev = torch.device("cuda", self.dev)
        mine = torch.tensor(self.sizes, dtype=torch.int64, device=dev)
        allsz = torch.empty(self.world * self.nchunks, dtype=torch.int64, device=dev)
        if self.world > 1:
            dist.all_gather_into_tensor(allsz, mine, group=self.group)      # ... and once everybody says so, the batch is complete
        else:
            allsz.copy_(mine)
        self.live = []
        self.bytes_moved = 0
        if self.rank != self.dst:
            return None, None
        sz = allsz.cpu().view(self.world, self.nchunks)
        counts = torch.as_tensor(_DevArray(self.base, (self.npat,)), device=dev)
        segs = []
        for r in range(self.world):
            off = 0
            for c in range(self.nchunks):
                n = int(sz[r, c])
                ptr = self.base + self.tup_off + 20 * (r * self.cap + off)
                segs.append(torch.as_tensor(_DevArray(ptr, (n, 5)), device=dev) if n else torch.zeros((0, 5), dtype=torch.int32, device=dev))
                off += n
                if r != self.dst:
                    self.bytes_moved += 20 * n
        self.bytes_moved += 4 * (self.npat - (self.e - self.b))
        return counts, segs

    def close(self):
        if self.rank == self.dst:
            self.eb.device_free(self.dev, self.base)
        else:
            self.eb.ipc_close(self.dev, self.base)


def _global_rank(r, group):
    return r if group is None else dist.get_global_rank(group, r)


class _DevArray:
    """Zero-copy view of library-owned device memory for torch (``torch.as_tensor(_DevArray(...), device=...)``)."""

    def __init__(self, ptr, shape, typestr="<i4"):
        self.__cuda_array_interface__ = {"shape": tuple(shape), "typestr": typestr, "data": (int(ptr), False), "version": 3}


def device_result_tensors(res, device):
    """DeviceResult -> (counts int32 [npat], tuples int32 [nocc, 5]) views over the library's HBM buffers."""
    c_ptr, o_ptr = res.device_ptrs()
    dev = torch.device("cuda", device)
    counts = torch.as_tensor(_DevArray(c_ptr, (res.npat,)), device=dev) if res.npat else torch.zeros(0, dtype=torch.int32, device=dev)
    occ = torch.as_tensor(_DevArray(o_ptr, (res.nocc, 5)), device=dev) if res.nocc else torch.zeros((0, 5), dtype=torch.int32, device=dev)
    return counts, occ


class ShardedSearcher:
    """Index replicated per GPU + pattern sharding + NCCL result gather.  `index` is an eds-bwt_b200 Index on this rank's GPU."""

    def __init__(self, index, group=None):
        self.index = index
        self.group = group
        self.world = dist.get_world_size(group)
        self.rank = dist.get_rank(group)

    def search(self, data, off, locate=True, lf_variant=0, dst=0):
        """data/off: the WHOLE batch as packed host arrays (uint8 bytes, uint64 offsets[npat+1]), identical on every rank.
        -> (counts uint32[npat] on every rank, structured occurrence array on rank dst / None elsewhere)"""
        from . import OCC_DTYPE       # noqa: the package name carries a hyphen; relative import works
        npat = len(off) - 1
        b, e = shard_range(npat, self.world, self.rank)
        dev = torch.device("cuda", self.index.device)
        lo, hi = int(off[b]), int(off[e])
        d_data = torch.from_numpy(np.ascontiguousarray(data[lo:hi]) if hi > lo else np.zeros(1, np.uint8)).to(dev)
        d_off = torch.from_numpy((off[b:e + 1] - off[b]).astype(np.int64)).to(dev)
        max_len = int((off[b + 1:e + 1] - off[b:e]).max()) if e > b else 0
        res = self.index.search_device(d_data.data_ptr(), d_off.data_ptr(), e - b, hi - lo, max_len, locate=locate,
                                       lf_variant=lf_variant, pat_base=b)
        try:
            c_loc, o_loc = device_result_tensors(res, self.index.device)
            counts, occ, _ = gather_results(c_loc, o_loc, npat, self.group, dst)
            counts_np = counts.cpu().numpy().view(np.uint32)
            occ_np = None
            if occ is not None:
                occ_np = occ.cpu().numpy().view(np.uint32).reshape(-1).view(OCC_DTYPE) if occ.numel() else np.zeros(0, OCC_DTYPE)
        finally:
            res.free()
        return counts_np, occ_np
