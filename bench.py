#!/usr/bin/env python
"""bench.py -- headline benchmark of the EDS-BWT batched pattern search on B200 (BASELINE.json metric:
"patterns/sec located (len 32)").

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

A "step" = one pass of the hot path (seed lookup + backward search + locate) over one batch of synthetic
patterns.  Workload at every N: cfg4 of SURVEY.md 8(d) -- the 250 Mbp synthetic chromosome-scale EDS
(index replicated on every GPU) with 10 M length-32 patterns PER GPU (weak scaling, patterns sharded, no
data-path collective).  value = patterns of all ranks / max-over-ranks device time.

  value     patterns already packed in HBM, result tuples left in HBM (edsbwt_search_device)
  e2e       the same batch through edsbwt_search with HOST buffers: H2D of the patterns and D2H of the
            counts + occurrence tuples inside the timed region
  roofline  dominant kernel against measured HBM bandwidth (see DESIGN.md "Measurement")
  cpu_baseline   the reference's own CPU search (oracle/_ref, unmodified sources) on a bounded sample

--impl reference times the reference's CPU implementation (all host cores) on the same index and prints
the same JSON line shape.
"""
import argparse
import ctypes
import glob
import importlib
import json
import os
import shutil
import subprocess
import sys
import tempfile
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "patterns/sec located (len 32)"
UNIT = "patterns/s"


def log(*a):
    print("[bench]", *a, file=sys.stderr, flush=True)


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--mbp", type=float, default=float(os.environ.get("EDSBWT_BENCH_MBP", "250")))
    ap.add_argument("--npat", type=int, default=int(os.environ.get("EDSBWT_BENCH_NPAT", "10000000")))
    ap.add_argument("--len", type=int, default=32, dest="plen")
    ap.add_argument("--seed-k", type=int, default=-1)
    ap.add_argument("--fused-cap", type=int, default=0)
    ap.add_argument("--fused-cap2", type=int, default=0)
    ap.add_argument("--workdir", default=os.environ.get("EDSBWT_BENCH_DIR", ""))
    ap.add_argument("--cpu-seconds", type=float, default=20.0, help="CPU budget of the cpu_baseline sample")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--keep", action="store_true")
    ap.add_argument("--chunk", type=int, default=0, help="patterns per pipeline chunk of the host-buffer search (0 = library default)")
    ap.add_argument("--config", default="cfg4", choices=["cfg4", "cfg5"],
                    help="cfg4: the headline (10 M length-32 patterns per GPU, weak scaling); cfg5: BASELINE.json's length sweep -- --total patterns of "
                         "length --len sharded over the GPUs (strong scaling), count-only and locate, on the same 250 Mbp index")
    ap.add_argument("--total", type=int, default=100000000, help="cfg5: patterns of the whole job")
    ap.add_argument("--gather", default="peer", choices=["peer", "nccl"], help="multi-GPU result gather: peer copies into an IPC arena on rank 0 (default) or ncclSend/ncclRecv")
    ap.add_argument("--gather-chunks", type=int, default=0, help="chunks per shard of the multi-GPU leg (gather of chunk c overlaps the search of chunk c+1); "
                    "0 = 2 chunks at 2 GPUs, 4 beyond (measured: profiles/r02_bench_n2_peer_*.json, r02_bench_n4.json, r02_bench_n8.json)")
    ap.add_argument("--engine", default="auto", choices=["auto", "items", "step"],
                    help="first-pass kernel: auto = the library default (step-loop kernel for bytes, item kernel for 2-bit words), or force one (A/B)")
    return ap.parse_args()


# ----------------------------------------------------------------------------------------------------
# clocks
# ----------------------------------------------------------------------------------------------------
class ClockSampler(threading.Thread):
    """Samples SM clocks and throttle reasons while the timed region runs: NVML every 2 ms (the timed region of the default
    run lasts tens of milliseconds), nvidia-smi as the fallback."""

    def __init__(self, device):
        super().__init__(daemon=True)
        self.device = device
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._halt = threading.Event()
        self.how = "nvml"
        try:
            import pynvml
            pynvml.nvmlInit()
            self._nv = pynvml
            self._h = pynvml.nvmlDeviceGetHandleByIndex(device)
            self.max_mhz = float(pynvml.nvmlDeviceGetMaxClockInfo(self._h, pynvml.NVML_CLOCK_SM))
        except Exception:
            self._nv = None
            self.how = "nvidia-smi"

    def _nvml_once(self):
        nv = self._nv
        self.samples.append(float(nv.nvmlDeviceGetClockInfo(self._h, nv.NVML_CLOCK_SM)))
        try:
            r = nv.nvmlDeviceGetCurrentClocksEventReasons(self._h)
        except Exception:
            r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self._h)
        for nm, bit in (("hw_slowdown", 0x8), ("hw_thermal_slowdown", 0x40), ("sw_thermal_slowdown", 0x20), ("sw_power_cap", 0x4)):
            if r & bit:
                self.reasons.add(nm)

    def _smi_once(self):
        q = "clocks.sm,clocks.max.sm,clocks_throttle_reasons.hw_slowdown,clocks_throttle_reasons.hw_thermal_slowdown," \
            "clocks_throttle_reasons.sw_thermal_slowdown,clocks_throttle_reasons.sw_power_cap"
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        out = subprocess.run(["nvidia-smi", "-i", str(self.device), "--query-gpu=" + q, "--format=csv,noheader,nounits"],
                             capture_output=True, text=True, timeout=5).stdout.strip().split(",")
        self.samples.append(float(out[0]))
        self.max_mhz = float(out[1])
        for nm, v in zip(names, out[2:]):
            if v.strip().lower() == "active":
                self.reasons.add(nm)

    def run(self):
        while not self._halt.is_set():
            try:
                if self._nv is not None:
                    self._nvml_once()
                else:
                    self._smi_once()
            except Exception:
                pass
            self._halt.wait(0.002 if self._nv is not None else 0.05)

    def stop(self):
        self._halt.set()
        self.join(timeout=6)
        med = float(np.median(self.samples)) if self.samples else None
        return {"sm_mhz": med, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons), "samples": len(self.samples), "how": self.how}


# ----------------------------------------------------------------------------------------------------
# workload
# ----------------------------------------------------------------------------------------------------
def workload_name(args):
    return "cfg4: synthetic %g Mbp chromosome-scale EDS (1 variant site / 100 bp, 2-8 alleles), %d length-%d patterns per GPU, locate" % (
        args.mbp, args.npat, args.plen)


NCU_CAPTURE = {"fused_search_kernel": "profiles/r02_fused_ncu_full.txt",       # ncu --set full captures at the default workload (cfg4, k = 14)
               "search_items_kernel": "profiles/r02_items_ncu_full.txt"}


def ncu_traffic(args, seed_k, kernel):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of the dominant kernel, from the committed ncu capture of
    this same command (only valid for the default workload the capture was taken on)."""
    if args.mbp != 250 or args.npat != 10000000 or args.plen != 32 or seed_k != 14 or kernel not in NCU_CAPTURE:
        return None, None
    try:
        tot = 0.0
        with open(os.path.join(ROOT, NCU_CAPTURE[kernel])) as f:
            for line in f:
                t = line.split()
                if len(t) >= 3 and t[0] in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
                    tot += float(t[1]) * {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0}[t[2]]
        return (tot or None), NCU_CAPTURE[kernel]
    except Exception:
        return None, None


def build_index_files(eb, synth, args, base):
    t0 = time.time()
    col = synth.chromosome_eds(args.mbp, seed=4)
    t1 = time.time()
    eb.build_from_collection(col.letters, col.str_off, col.seg_first, base, device=0)
    t2 = time.time()
    log("EDS generated in %.1fs (n=%d rows, %d strings, %d segments); index files built on the GPU + written in %.1fs" % (
        t1 - t0, col.n_rows, col.nstr, col.n_seg, t2 - t1))
    return col


def cpu_reference_run(base, pats, seconds, cores, note, per=None):
    """Times the reference's own CPU search on `cores` independent processes (it is single-threaded:
    mainEDS-BWT.cpp:66), each on its own shard of a bounded sample.  -> dict for cpu_baseline."""
    from oracle import run as R
    exe = R.ref_binary("rb1")
    kind = "reference"
    if exe is None:
        if not os.path.exists(R.PORT):
            R.build()
        exe, kind = R.PORT, "port"
    d = tempfile.mkdtemp(prefix="edsbwt_cpu_")
    try:
        def run_shards(per):
            procs, files = [], []
            for c in range(cores):
                pf = os.path.join(d, "p%d.txt" % c)
                with open(pf, "wb") as f:
                    for r in pats[c * per:(c + 1) * per]:
                        f.write(r.tobytes() + b"\n")
                files.append(pf)
            t0 = time.time()
            for pf in files:
                cmd = [exe, base, pf] if kind == "reference" else [exe, base, pf, "--locate", "1", "--csv", pf + "output.csv"]
                procs.append(subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True))
            took, found, rows = [], [], []
            for c, p in enumerate(procs):
                out = p.communicate()[0]
                if kind == "reference":
                    import re
                    m = re.search(r"bs took:([0-9.eE+-]+)", out)
                    took.append(float(m.group(1)) if m else float("nan"))
                    found += [int(x) for x in re.findall(r"Found (\d+) starting", out)]
                else:
                    took.append(json.loads(out.strip().splitlines()[-1])["seconds"])
                # the reference's own csv rows (<patterns>output.csv, EDSBWTsearch.cpp:1310-1332), pattern numbers made sample-global
                csv = files[c] + "output.csv"
                if os.path.exists(csv):
                    with open(csv) as f:
                        f.readline()
                        for line in f:
                            t = line.split()
                            if len(t) == 5:
                                rows.append((int(t[0]) - 1 + c * per, int(t[1]), int(t[2]), int(t[3]), int(t[4])))
            return time.time() - t0, max(took), found, rows
        # calibrate with one pattern per core, then size the sample for ~`seconds` of search time
        if per is None:
            wall1, took1, _, _ = run_shards(1)
            per = int(max(1, min(len(pats) // cores, seconds / max(took1, 1e-3))))
        wall, took, found, rows = run_shards(per)
        val = per * cores / max(took, 1e-9)
        return {"value": val, "unit": UNIT, "cores": cores, "kind": kind,
                "sample": "%d patterns (%d per process x %d processes) of the same batch, %s; max 'bs took' %.2fs (index load excluded, wall %.1fs)" % (
                    per * cores, per, cores, note, took, wall),
                "search_seconds": took, "counts": found, "rows": rows, "n": per * cores, "per": per}
    finally:
        shutil.rmtree(d, ignore_errors=True)


_ORIG_AFFINITY = None


def bind_to_gpu_numa_node(device):
    """Pin this process to the CPUs next to its GPU (NVML's ideal affinity) before any pinned buffer is allocated, so that
    the host side of the e2e copies lives on the GPU's own NUMA node / PCIe root (matters with 8 ranks on one box)."""
    global _ORIG_AFFINITY
    try:
        import pynvml
        _ORIG_AFFINITY = os.sched_getaffinity(0)
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(device)
        words = (os.cpu_count() + 63) // 64
        mask = pynvml.nvmlDeviceGetCpuAffinity(h, words)
        cpus = [64 * w + b for w in range(words) for b in range(64) if (mask[w] >> b) & 1]
        if cpus:
            os.sched_setaffinity(0, set(cpus) & os.sched_getaffinity(0) or set(cpus))
            log("device %d: bound to %d CPUs (%d..%d)" % (device, len(cpus), cpus[0], cpus[-1]))
    except Exception as e:      # noqa
        log("no NUMA binding:", e)


# ----------------------------------------------------------------------------------------------------
_REAL_STDOUT = None


def emit(obj):
    """The ONE JSON line of the contract, on the process's real stdout."""
    line = (json.dumps(obj) + "\n").encode()
    if _REAL_STDOUT is None:
        sys.stdout.write(line.decode()); sys.stdout.flush()
    else:
        os.write(_REAL_STDOUT, line)


def main():
    global _REAL_STDOUT
    args = parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if world == 1 and args.gpus > 1 and args.impl == "ours":
        # convenience: relaunch under torchrun
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(args.gpus),
               "--master-addr", "127.0.0.1", "--master-port", "29541", os.path.abspath(__file__)] + sys.argv[1:]
        sys.exit(subprocess.call(cmd))
    if args.impl == "reference" and rank != 0:
        return 0
    # libraries (NCCL's version banner, torchrun notices) write to fd 1: keep stdout for the JSON line only
    sys.stdout.flush()
    _REAL_STDOUT = os.dup(1)
    os.dup2(2, 1)

    import torch
    import torch.distributed as dist
    eb = importlib.import_module("eds-bwt_b200")
    synth = importlib.import_module("eds-bwt_b200.synth")

    workdir = args.workdir or os.path.join("/dev/shm" if os.path.isdir("/dev/shm") else tempfile.gettempdir(), "edsbwt_bench_%d" % os.getuid())
    base = os.path.join(workdir, "cfg4_%g" % args.mbp)
    done_flag = base + ".done"

    if args.impl == "reference":
        return reference_arm(args, eb, synth, workdir, base, done_flag)

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- the product has no CPU path (use --impl reference for the CPU baseline)")
    torch.cuda.set_device(local_rank)
    bind_to_gpu_numa_node(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    eb.lib()

    # ---- index files: rank 0 generates + builds (GPU suffix sort), everybody opens
    if rank == 0:
        os.makedirs(workdir, exist_ok=True)
        if not os.path.exists(done_flag):
            build_index_files(eb, synth, args, base)
            open(done_flag, "w").close()
    if world > 1:
        dist.barrier()
    t0 = time.time()
    idx = eb.Index(base, device=local_rank, seed_k=args.seed_k, verify=False, fused_cap=args.fused_cap, fused_cap2=args.fused_cap2,
                   engine={"auto": eb.ENGINE_FUSED, "items": eb.ENGINE_ITEMS, "step": eb.ENGINE_FUSED_STEP}[args.engine])
    info = idx.info()
    log("rank %d: index open in %.1fs: n=%d nText=%d nSeg=%d seed_k=%d seed_intervals=%d device_bytes=%.2f GB" % (
        rank, time.time() - t0, info.n, info.n_text, info.n_seg, info.seed_k, info.seed_intervals, info.device_bytes / 1e9))

    if args.config == "cfg5":
        out = cfg5_leg(args, eb, synth, idx, info, base, rank, world, local_rank)
        idx.close()
        if world > 1:
            dist.barrier()
            dist.destroy_process_group()
        if rank == 0:
            emit(out)
        return 0

    # ---- patterns of this rank (every rank regenerates the collection; seeds differ per rank)
    t0 = time.time()
    col = synth.chromosome_eds(args.mbp, seed=4)
    pats = synth.sample_patterns(col, args.npat, args.plen, seed=44 + rank)
    del col
    data, off = synth.pack_fixed(pats)
    log("rank %d: %d patterns sampled in %.1fs" % (rank, args.npat, time.time() - t0))
    h_data = torch.from_numpy(data).pin_memory()
    h_off = torch.from_numpy(off.view(np.int64)).pin_memory()
    d_data = h_data.cuda(non_blocking=False)
    d_off = h_off.cuda(non_blocking=False)
    total_bytes = int(data.size)
    torch.cuda.synchronize()

    def step_device():
        r = idx.search_device(d_data.data_ptr(), d_off.data_ptr(), args.npat, total_bytes, args.plen, locate=True)
        st = idx.last_stats()
        nocc = r.nocc
        r.free()
        return st, nocc

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # counters pass (untimed): canonical byte model I,S,E,L,O + Occ blocks actually fetched
    idx.set_stats(True, 12)
    st_c, nocc = step_device()
    # ... and the counters of the configuration that is timed (its own seed depth): Occ blocks / link records / seed
    # intervals actually fetched = the algorithmic bytes of roofline.achieved
    idx.set_stats(True, max(12, int(info.seed_k)))
    st_a, _ = step_device()
    idx.set_stats(False, 12)
    for _ in range(max(args.warmup, 3) - 1):
        step_device()

    # ---- timed region: `value`
    barrier()
    sampler = ClockSampler(local_rank)
    sampler.start()
    dev_ms, lf_ms, loc_ms, launches = 0.0, 0.0, 0.0, 0
    t0 = time.perf_counter()
    torch.cuda.nvtx.range_push("timed")
    for _ in range(args.steps):
        st, _ = step_device()
        dev_ms += st["total_ms"]; lf_ms += st["search_kernel_ms"]; loc_ms += st["locate_kernel_ms"]; launches += st["kernel_launches"]
    torch.cuda.nvtx.range_pop()
    barrier()
    wall_ms = (time.perf_counter() - t0) * 1e3
    clocks = sampler.stop()

    # ---- the other first-pass kernel on the same device-resident batch (the item kernel takes 2-bit words: for bytes that adds the
    #      packing pass, which is why the library gives bytes to the step-loop kernel): reported next to the roofline of `value`
    other = None
    if args.engine == "auto":
        idx.set_engine(eb.ENGINE_ITEMS)
        o_ms = o_k = 0.0
        for it in range(2 + args.steps):
            st, _ = step_device()
            if it >= 2:
                o_ms += st["total_ms"]; o_k += st["search_kernel_ms"]
        idx.set_engine(eb.ENGINE_FUSED)
        other = {"kernel": "search_items_kernel", "ms_per_step": o_ms / args.steps, "kernel_ms_per_step": o_k / args.steps,
                 "value": args.npat / (o_ms / args.steps * 1e-3), "note": "engine = EDSBWT_ENGINE_ITEMS on the same batch of bytes: 2-bit packing pass + item kernel + locate"}

    # ---- e2e: host buffers through edsbwt_search_into (pinned caller buffers; H2D + D2H inside the timed region)
    p_data = eb.pinned_empty(data.size, np.uint8); p_data[:] = data
    p_off = eb.pinned_empty(off.size, np.uint64); p_off[:] = off
    p_counts = eb.pinned_empty(args.npat, np.uint32)
    p_occ = eb.pinned_empty(int(nocc) + 1024, eb.OCC_DTYPE)
    if args.chunk:
        idx.set_chunk(args.chunk)
    # (a) the generic call: pattern bytes + uint64 offsets (variable-length batches, the reference's pattern file)
    e2o_ms = 0.0
    for it in range(2 + args.steps):
        barrier()
        t1 = time.perf_counter()
        got = idx.search_into(p_data, p_off, p_counts, p_occ, locate=True)
        dt = (time.perf_counter() - t1) * 1e3
        if it >= 2:
            e2o_ms += dt
    counts_generic = np.array(p_counts, copy=True)
    # (b) the call for equal-length batches of bytes: no offset array crosses PCIe
    e2f_ms = 0.0
    for it in range(2 + args.steps):
        barrier()
        t1 = time.perf_counter()
        got = idx.search_into_fixed(p_data, args.plen, p_counts, p_occ, locate=True)
        dt = (time.perf_counter() - t1) * 1e3
        if it >= 2:
            e2f_ms += dt
    assert np.array_equal(counts_generic, np.asarray(p_counts)), "fixed-length and offset-array entry points disagree"
    counts_fixed = np.array(p_counts, copy=True)
    # (c) the headline e2e: the caller holds the batch as 2-bit words (edsbwt_search_into_packed): a quarter of the bytes over PCIe.
    #     The host packer that makes them from bytes is timed separately (it belongs to reading the pattern file, not to a search)
    stride = (args.plen + 31) // 32
    p_words = eb.pinned_empty(args.npat * stride, np.uint64)
    p_inval = eb.pinned_empty((args.npat + 31) // 32, np.uint32)
    t1 = time.perf_counter()
    _, _, nbad = idx.pack_patterns(p_data, args.plen, nthreads=os.cpu_count() or 1, words=p_words, invalid=p_inval)
    pack_ms = (time.perf_counter() - t1) * 1e3
    e2e_ms, d2h = 0.0, 0
    for it in range(2 + args.steps):
        barrier()
        t1 = time.perf_counter()
        got = idx.search_into_packed(p_words, p_inval, args.plen, args.npat, p_counts, p_occ, locate=True)
        dt = (time.perf_counter() - t1) * 1e3
        if it >= 2:
            e2e_ms += dt
            d2h = p_counts.nbytes + got * eb.OCC_DTYPE.itemsize
    h2d = p_words.nbytes + p_inval.nbytes
    assert nbad == 0 and np.array_equal(counts_fixed, np.asarray(p_counts)), "packed and byte entry points disagree"
    gpu_counts = np.array(p_counts, copy=True)
    gpu_occ_total = int(got)
    gpu_occ_head = np.array(p_occ[:min(int(got), 1 << 18)], copy=True)      # tuples of the first patterns: row parity against the reference sample
    # ---- multi-GPU result gather (north star: "only an NCCL/NVLink gather of per-pattern counts and occurrence buffers"):
    #      shard search + all-gather of the counts + gather-v of the tuples to rank 0, timed on the device
    # fewer chunks = fewer kernel tails, more chunks = less of the last transfer exposed: at 2 GPUs rank 0 receives 102 MB and
    # 2 chunks win (3.81 vs 4.06 ms), with 3 or 7 senders the inbound link is the longer pole and 4 chunks hide it
    gather_ms, gather_bytes, gather_parity, nchunks = 0.0, 0, None, (args.gather_chunks or (2 if world == 2 else 4))
    if world > 1:
        # every shard is searched in chunks; the counts and tuples of chunk c are put on the wire (ncclSend / ncclRecv, one group
        # per chunk) as soon as it is searched and travel while chunk c + 1 is searched.  Timed on the device: CUDA events on this
        # rank's stream from before the first chunk to after the last transfer has landed, max over ranks below.
        sh = importlib.import_module("eds-bwt_b200.sharded")
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        gnpat = args.npat * world
        d_off64 = d_off
        # default: the gather over peer memory (CUDA IPC arena on rank 0, asynchronous NVLink peer copies, no size exchange);
        # --gather nccl: ncclSend / ncclRecv per chunk with the sizes exchanged over a gloo group (A/B)
        peer = sh.PeerGather(eb, idx, gnpat, nchunks, int(nocc * 1.25) + 4096, None, 0) if args.gather == "peer" else None
        size_group = dist.new_group(backend="gloo") if peer is None else None
        for it in range(2 + args.steps):
            barrier()
            ev0.record()
            if peer is not None:
                peer.begin()
            g = sh.ChunkedGather(gnpat, nchunks, None, size_group, 0) if peer is None else peer
            live = []
            for c in range(nchunks):
                cb, ce = sh.chunk_range(0, args.npat, nchunks, c)
                # (offsets are absolute: the byte pointer stays the batch's, the offset pointer moves to the chunk)
                r = idx.search_device(d_data.data_ptr(), d_off64.data_ptr() + cb * 8, ce - cb, (ce - cb) * args.plen, args.plen,
                                      locate=True, pat_base=rank * args.npat + cb)
                live.append(r)
                if peer is not None:
                    peer.push(c, r)
                else:
                    c_loc, o_loc = sh.device_result_tensors(r, local_rank)
                    g.post(c, c_loc, o_loc)
            allc, segs = g.finish()
            ev1.record()
            torch.cuda.synchronize()
            if it >= 2:
                gather_ms += ev0.elapsed_time(ev1)
                gather_bytes = g.bytes_moved
            if it == 2 + args.steps - 1:
                # gather parity: what arrived on rank 0 against what every rank computed (sums over its own shard, all-reduced)
                loc = torch.zeros(3, dtype=torch.int64, device="cuda")
                for r in live:
                    c_loc, o_loc = sh.device_result_tensors(r, local_rank)
                    loc[0] += c_loc.to(torch.int64).sum(); loc[1] += o_loc.shape[0]
                    if o_loc.shape[0]:
                        loc[2] += (o_loc[:, 0].to(torch.int64) & 0xffffffff).sum()
                dist.all_reduce(loc)
                if rank == 0:
                    got3 = [int(allc.to(torch.int64).sum()), sum(int(x.shape[0]) for x in segs),
                            sum(int((x[:, 0].to(torch.int64) & 0xffffffff).sum()) for x in segs if x.shape[0])]
                    asc = all(bool((x[1:, 0].to(torch.int64) >= x[:-1, 0].to(torch.int64)).all()) for x in segs if x.shape[0] > 1)
                    own = bool(torch.equal(allc[:args.npat].cpu(), torch.from_numpy(gpu_counts.view(np.int32))))
                    gather_parity = "ok (%d counts, %d tuples: sums of counts / tuple count / pattern ids match all ranks; segments ascending)" % (
                        gnpat, got3[1]) if (got3 == [int(v) for v in loc.tolist()] and asc and own) else "MISMATCH %r vs %r" % (got3, loc.tolist())
            for r in live:
                r.free()
            del allc, segs, g
        if peer is not None:
            barrier()
            peer.close()

    def allmax(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    dev_ms_max = allmax(dev_ms)
    wall_ms_max = allmax(wall_ms)
    e2e_ms_max = allmax(e2e_ms)
    e2o_ms_max = allmax(e2o_ms)
    gather_ms_max = allmax(gather_ms) if world > 1 else 0.0
    e2f_ms_max = allmax(e2f_ms)
    total_pat = args.npat * world * args.steps
    value_no_gather = total_pat / (dev_ms_max * 1e-3)
    # N > 1: the north star's collective is part of the job, so it is part of `value` (shard search + NCCL gather of counts and
    # tuples on rank 0, overlapped); N = 1 has nothing to gather
    value = total_pat / (gather_ms_max * 1e-3) if world > 1 else value_no_gather
    step_ms_max = gather_ms_max if world > 1 else dev_ms_max
    e2e_val = total_pat / (e2e_ms_max * 1e-3)

    out = None
    if rank == 0:
        peaks = {}
        try:
            with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
                peaks = json.load(f)
        except Exception:
            pass
        peak = float(peaks.get("hbm_gbs", 6553.6))
        peak_src = "MEASURED_PEAKS.json hbm_gbs" if "hbm_gbs" in peaks else "fallback 6553.6 GB/s (B200_PROFILING.md)"
        # dominant kernel: the search kernel.  Algorithmic bytes = 64 B per Occ block fetched + 8 B per link
        # record + 8 B per seed interval read + the pattern itself: 2-bit words and one 8-byte offset per pattern for the item
        # kernel, the raw bytes and two offsets for the step-loop kernel.  Canonical SURVEY model printed next to it.
        occ_blocks = st_a["occ_blocks"]
        kernel = "search_items_kernel" if args.engine == "items" else "fused_search_kernel"
        pat_bytes = (8.0 * ((args.plen + 31) // 32) + 8.0) if args.engine == "items" else (args.plen + 16.0)
        alg_bytes = 64.0 * occ_blocks + 8.0 * st_a["E"] + pat_bytes * args.npat + 8.0 * st_a["I"]
        lf_ms_step = lf_ms / args.steps
        achieved = alg_bytes / (lf_ms_step * 1e-3) / 1e9 if lf_ms_step > 0 else 0.0
        traffic, traffic_src = ncu_traffic(args, int(info.seed_k), kernel)
        canon = 8.0 * (args.npat + st_c["I"]) + 128.0 * st_c["S"] + 16.0 * st_c["E"] + 64.0 * st_c["L"] + 28.0 * st_c["O"]
        # random-access ceilings measured in this run (csrc/gather_bench.cu: the kernel's own 64-byte fetch, one thread per block):
        # best of a few independent-access configurations over a buffer of the size of the Occ array, over 4 GiB, and for the
        # kernel's own mix (block fetches over the Occ-sized buffer + 18 % 16-byte gathers over 5 GB)
        occ_bytes = max(64 << 20, int(info.n // 128 + 1) * 64)
        try:
            g_idx = max(eb.gather_bench_cfg(local_rank, occ_bytes, 1 << 27, 2, 64, ilp, 0, False)[0] for ilp in (4, 8))
            g_big = max(eb.gather_bench_cfg(local_rank, 4 << 30, 1 << 27, 2, 64, ilp, 0, False)[0] for ilp in (4, 8))
            g_mix = max(eb.gather_bench_cfg(local_rank, occ_bytes, 1 << 27, 2, 64, ilp, 0, False, 5 << 30, 0.18)[0] for ilp in (4, 8))
            g_dep = eb.gather_bench_cfg(local_rank, occ_bytes, 1 << 27, 2, 64, 1, 0, True)[0]
        except Exception as e:      # noqa
            log("gather bench failed:", e)
            g_big = g_idx = g_mix = g_dep = None
        out = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
            "ms_per_step": step_ms_max / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "u32", "data": "synthetic",
            "config": {"workload": workload_name(args), "index_rows": info.n, "strings": info.n_text, "segments": info.n_seg,
                       "patterns_per_gpu": args.npat, "pattern_len": args.plen, "seed_k": info.seed_k, "locate": "sa-table",
                       "search_kernel": "search_items_kernel (persistent warps, lane refill)" if args.engine == "items" else "fused_search_kernel (step loop, one thread per pattern; the library's choice for batches of bytes)",
                       "parallelism": "patterns sharded x%d, index replicated" % world,
                       "l2": "inputs larger than L2 (patterns %.0f MB + index %.2f GB per step; no flush needed)" % (data.nbytes / 1e6, info.device_bytes / 1e9)},
            "wall_ms_per_step": wall_ms_max / args.steps,
            "occurrences_per_step": int(nocc),
            "value_no_gather": value_no_gather, "ms_per_step_no_gather": dev_ms_max / args.steps,
            "e2e": {"value": e2e_val, "unit": UNIT, "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
                    "ms_per_step": e2e_ms_max / args.steps,
                    "call": "edsbwt_search_into_packed (pinned host buffers; patterns as 2-bit words, 8 B per length-32 pattern, + the invalid bitmap)",
                    "host_packer": {"call": "edsbwt_pack_patterns", "ms": pack_ms, "threads": os.cpu_count() or 1, "patterns_per_s": args.npat / (pack_ms * 1e-3),
                                    "note": "bytes -> words on the host, once per pattern set; not inside the e2e region"},
                    "ascii_fixed": {"value": total_pat / (e2f_ms_max * 1e-3), "ms_per_step": e2f_ms_max / args.steps, "h2d_bytes_per_step": int(data.nbytes),
                                    "call": "edsbwt_search_into_fixed (raw pattern bytes, packed on the device)"},
                    "with_offset_array": {"value": total_pat / (e2o_ms_max * 1e-3), "ms_per_step": e2o_ms_max / args.steps,
                                          "h2d_bytes_per_step": int(data.nbytes + off.nbytes), "call": "edsbwt_search_into"}},
            "gather": ({"in_value": True, "chunks_per_shard": nchunks, "ms_per_step": gather_ms_max / args.steps, "nvlink_bytes_per_step": int(gather_bytes),
                        "GBps_into_rank0": gather_bytes / (gather_ms_max / args.steps * 1e-3) / 1e9 if gather_ms_max else None,
                        "gather_parity": gather_parity,
                        "transport": "NVLink peer copies into a CUDA-IPC arena on rank 0 (edsbwt_result_push)" if args.gather == "peer" else "ncclSend / ncclRecv, sizes over gloo",
                        "what": "per chunk: search, then its counts and tuples go to rank 0 while the next chunk is searched; one all-gather of the chunk sizes per step; CUDA events, max over ranks"}
                       if world > 1 else None),
            "gpu_launches": int(launches),
            "clocks": clocks,
            "roofline": {"bound": "hbm", "kernel": kernel, "achieved": achieved, "peak": peak, "unit": "GB/s",
                         "frac": achieved / peak if peak else None, "traffic": traffic, "traffic_source": traffic_src, "peak_source": peak_src,
                         "alg_bytes_per_step": alg_bytes, "kernel_ms_per_step": lf_ms_step,
                         "kernel_share_of_step": lf_ms_step / (dev_ms / args.steps) if dev_ms else None,
                         "item_kernel_on_the_same_batch": (dict(other, achieved=(64.0 * occ_blocks + 8.0 * st_a["E"] + 16.0 * args.npat + 8.0 * st_a["I"]) / (other["kernel_ms_per_step"] * 1e-3) / 1e9,
                                                                 frac=(64.0 * occ_blocks + 8.0 * st_a["E"] + 16.0 * args.npat + 8.0 * st_a["I"]) / (other["kernel_ms_per_step"] * 1e-3) / 1e9 / peak) if other else None),
                         "gather64_gbs_4GiB": g_big, "gather64_gbs_occ_sized": g_idx, "gather_mix_gbs": g_mix, "gather64_dependent_chains_gbs": g_dep,
                         "frac_of_gather64": (achieved / g_idx) if g_idx else None, "frac_of_gather_mix": (achieved / g_mix) if g_mix else None,
                         "gather_note": "independent 64-byte block fetches, best of ilp 4 / 8 at full occupancy; the Occ array is what the kernel gathers from",
                         "locate_kernel_ms_per_step": loc_ms / args.steps, "overflow_patterns": st_a["overflow_patterns"], "occ_blocks": occ_blocks,
                         "lf_intervals": st_a["lf_intervals"], "link_records": st_a["E"], "seed_intervals_read": st_a["I"],
                         "canonical_model": {"k": 12, "I": st_c["I"], "S": st_c["S"], "E": st_c["E"], "L": st_c["L"], "O": st_c["O"],
                                             "bytes_per_step": canon, "bytes_per_pattern": canon / args.npat,
                                             "GBps_over_whole_step": canon / (dev_ms / args.steps * 1e-3) / 1e9}},
        }
    idx.close()
    if rank == 0 and world == 1 and not args.no_cpu:
        try:
            if _ORIG_AFFINITY:
                os.sched_setaffinity(0, _ORIG_AFFINITY)          # the CPU baseline gets every host core again
            cb = cpu_reference_run(base, pats, args.cpu_seconds, os.cpu_count() or 1, "locate (RECOVERBW=1)")
            ref_counts, ref_rows, nref = cb.pop("counts"), cb.pop("rows"), cb.pop("n"); cb.pop("per", None)
            if len(ref_counts) == nref:         # full-size parity check on the sample: the reference's own counts AND csv rows
                ngpu = int(gpu_counts[:nref].astype(np.int64).sum())
                gpu_rows = sorted(zip(gpu_occ_head["pat"][:ngpu].tolist(), gpu_occ_head["str"][:ngpu].tolist(), gpu_occ_head["seg"][:ngpu].tolist(),
                                      gpu_occ_head["idx_in_seg"][:ngpu].tolist(), gpu_occ_head["off"][:ngpu].tolist()))
                ok_c = ref_counts == gpu_counts[:nref].tolist()
                ok_r = ngpu <= len(gpu_occ_head) and sorted(ref_rows) == gpu_rows
                cb["parity_vs_gpu"] = ("counts+rows ok (%d patterns, %d rows)" % (nref, len(ref_rows))) if (ok_c and ok_r) else (
                    "MISMATCH (counts %s, rows %s)" % ("ok" if ok_c else "differ", "ok" if ok_r else "differ"))
            out["cpu_baseline"] = cb
        except Exception as e:
            out["cpu_baseline"] = {"value": None, "unit": UNIT, "cores": 0, "kind": "unavailable", "sample": "failed: %r" % (e,)}
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    if rank == 0:
        emit(out)
        if not args.keep and not os.environ.get("EDSBWT_BENCH_KEEP"):
            pass        # index files stay in the work dir for the reference arm that the driver runs next
    return 0


def device_patterns(col, npat, m, seed, device):
    """synth.sample_patterns' 90/10 recipe on the GPU (data preparation only, torch): 100 M patterns of length 256 are 25.6 GB,
    too slow to sample with numpy on the host.  -> uint8 tensor [npat, m] on `device`"""
    import torch
    g = torch.Generator(device=device); g.manual_seed(seed)
    rng = np.random.default_rng(seed)
    acgt = torch.tensor(list(b"ACGT"), dtype=torch.uint8, device=device)
    out = torch.empty((npat, m), dtype=torch.uint8, device=device)
    n_path = int(round(npat * 0.9))
    ar = torch.arange(m, device=device, dtype=torch.int64)
    H = torch.from_numpy(synth_haplotype(col, rng)).to(device)
    step = max(1, (1 << 28) // m)
    for b in range(0, npat, step):
        e = min(npat, b + step)
        np_rows = max(0, min(e, n_path) - b)
        if np_rows:
            st = torch.randint(0, H.numel() - m + 1, (np_rows,), generator=g, device=device)
            out[b:b + np_rows] = H[st[:, None] + ar[None, :]]
        if e - b > np_rows:
            out[b + np_rows:e] = acgt[torch.randint(0, 4, (e - b - np_rows, m), generator=g, device=device)]
    n_mut = int(round(n_path * 0.1))                      # one substitution in a tenth of the path patterns
    who = torch.randperm(n_path, generator=g, device=device)[:n_mut]
    where = torch.randint(0, m, (n_mut,), generator=g, device=device)
    cur = out[who, where]
    code = torch.searchsorted(acgt.to(torch.int16), cur.to(torch.int16))
    out[who, where] = acgt[(code + torch.randint(1, 4, (n_mut,), generator=g, device=device)) % 4]
    return out[torch.randperm(npat, generator=g, device=device)].contiguous()


def synth_haplotype(col, rng):
    import importlib
    return importlib.import_module("eds-bwt_b200.synth").haplotype(col, rng)


def cfg5_leg(args, eb, synth, idx, info, base, rank, world, local_rank):
    """BASELINE.json config 5: --total patterns of length --len on the cfg4 index, sharded over the ranks (strong scaling),
    count-only and locate; a sample of this rank's patterns is checked against the unmodified reference (counts and rows)."""
    import torch
    import torch.distributed as dist
    m = args.plen
    lo, hi = (args.total * rank) // world, (args.total * (rank + 1)) // world
    npat = hi - lo
    col = synth.chromosome_eds(args.mbp, seed=4)
    t0 = time.time()
    d_pats = device_patterns(col, npat, m, 55 + m + 1000 * rank, torch.device("cuda", local_rank))
    del col
    d_off = torch.arange(npat + 1, dtype=torch.int64, device=d_pats.device) * m
    torch.cuda.synchronize()
    log("rank %d: %d patterns of length %d sampled on the device in %.1fs" % (rank, npat, m, time.time() - t0))
    # locate: a length-8 key has thousands of occurrences -- locate only as many patterns as give ~2^27 tuples (cap stated in the line)
    r = idx.search_device(d_pats.data_ptr(), d_off.data_ptr(), min(npat, 1 << 16), min(npat, 1 << 16) * m, m, locate=False)
    c0, _ = r.fetch(); r.free()
    per = max(1.0, float(c0.astype(np.float64).mean()))
    nloc = int(min(npat, max(1024, (1 << 27) / per)))

    def run(n, locate):
        ms = 0.0; nocc = 0
        for it in range(max(args.warmup, 3) + args.steps):
            torch.cuda.synchronize()
            if world > 1:
                dist.barrier()
            rr = idx.search_device(d_pats.data_ptr(), d_off.data_ptr(), n, n * m, m, locate=locate)
            st = idx.last_stats(); nocc = rr.nocc; rr.free()
            if it >= max(args.warmup, 3):
                ms += st["total_ms"]
        return ms, nocc

    ms_count, _ = run(npat, False)
    ms_loc, nocc = run(nloc, True)

    def allmax(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    ms_count, ms_loc = allmax(ms_count), allmax(ms_loc)
    # oracle sample on rank 0: the reference binary on the first patterns of the shard
    parity = None
    if rank == 0 and not args.no_cpu:
        nref = 256 if m >= 16 else 64
        rr = idx.search_device(d_pats.data_ptr(), d_off.data_ptr(), nref, nref * m, m, locate=True)
        gc, go = rr.fetch(); rr.free()
        cb = cpu_reference_run(base, d_pats[:nref].cpu().numpy(), args.cpu_seconds, min(os.cpu_count() or 1, nref), "cfg5 m=%d" % m, per=max(1, nref // min(os.cpu_count() or 1, nref)))
        n = cb["n"]
        gsub = go[go["pat"] < n]
        rows = sorted(zip(gsub["pat"].tolist(), gsub["str"].tolist(), gsub["seg"].tolist(), gsub["idx_in_seg"].tolist(), gsub["off"].tolist()))
        ok = cb["counts"] == gc[:n].tolist() and sorted(cb["rows"]) == rows
        parity = {"kind": cb["kind"], "patterns": n, "rows": len(cb["rows"]), "result": "counts+rows ok" if ok else "MISMATCH",
                  "reference_patterns_per_s": cb["value"], "cores": cb["cores"]}
    if rank != 0:
        return None
    tot_count = args.total * args.steps
    return {"metric": "patterns/sec located (len %d)" % m, "value": (nloc * world * args.steps) / (ms_loc * 1e-3), "unit": UNIT, "n_gpus": world,
            "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": ms_loc / args.steps, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "u32", "data": "synthetic (patterns sampled on the device, 90/10 recipe, seed 55+m)",
            "config": {"workload": "cfg5: %d patterns of length %d on the synthetic %g Mbp EDS, sharded over %d GPU(s)" % (args.total, m, args.mbp, world),
                       "patterns_per_gpu": npat, "pattern_len": m, "seed_k": info.seed_k,
                       "located_patterns_per_gpu": nloc, "locate_cap": "locate leg runs on the first %d patterns of each shard (~2^27 tuples; a length-%d key has %.1f occurrences on average)" % (nloc, m, per)},
            "count_only": {"value": tot_count / (ms_count * 1e-3), "unit": UNIT, "ms_per_step": ms_count / args.steps, "patterns_per_step": args.total},
            "locate": {"value": (nloc * world * args.steps) / (ms_loc * 1e-3), "unit": UNIT, "ms_per_step": ms_loc / args.steps, "occurrences_per_step_rank0": int(nocc)},
            "parity_vs_reference": parity}


def reference_arm(args, eb, synth, workdir, base, done_flag):
    """--impl reference: the reference's CPU search, all host cores, same index / pattern recipe."""
    os.makedirs(workdir, exist_ok=True)
    if not os.path.exists(done_flag):
        import torch
        if not torch.cuda.is_available():
            emit({"impl": "reference", "unavailable": "index files for cfg4 are built by the GPU suffix sorter; no CUDA device here"})
            return 0
        build_index_files(eb, synth, args, base)
        open(done_flag, "w").close()
    col = synth.chromosome_eds(args.mbp, seed=4)
    pats = synth.sample_patterns(col, min(args.npat, 200000), args.plen, seed=44)
    del col
    cores = os.cpu_count() or 1
    vals, secs = [], []
    base_info = None
    # bounded: the whole --steps K --warmup W run gets ~150 s of search time; one calibration, then the same sample size
    budget = min(args.cpu_seconds, 150.0 / max(1, args.warmup + args.steps))
    per = None
    for it in range(args.warmup + args.steps):
        r = cpu_reference_run(base, pats[it * 4096:], budget, cores, "locate (RECOVERBW=1), step %d" % it, per=per)
        per = r["per"]; r.pop("rows", None)
        if it >= args.warmup:
            vals.append(r["value"]); secs.append(r["search_seconds"]); base_info = r
    v = float(np.mean(vals))
    out = {"impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
           "ms_per_step": float(np.mean(secs)) * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u32",
           "data": "synthetic", "config": {"workload": workload_name(args)},
           "cpu_baseline": {"value": v, "unit": UNIT, "cores": cores, "kind": base_info["kind"], "sample": base_info["sample"]},
           "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    emit(out)
    return 0


if __name__ == "__main__":
    sys.exit(main())
