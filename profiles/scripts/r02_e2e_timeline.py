"""Round 2: device timeline of ONE edsbwt_search_into_packed call at cfg4 (CUPTI through torch.profiler: kernels and copies of the
whole process, the library's included).  Prints every activity longer than 20 us with its stream, start and duration, and per-stream
busy time, so that the idle stretches of the chunk pipeline can be read off."""
import importlib
import json
import os
import sys

import numpy as np

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), "..", ".."))
sys.path.insert(0, ROOT)


def main():
    import torch
    from torch.profiler import profile, ProfilerActivity
    npat, plen = 10_000_000, 32
    eb = importlib.import_module("eds-bwt_b200")
    synth = importlib.import_module("eds-bwt_b200.synth")
    workdir = os.path.join("/dev/shm", "edsbwt_bench_%d" % os.getuid())
    os.makedirs(workdir, exist_ok=True)
    base = os.path.join(workdir, "cfg4_250")
    col = synth.chromosome_eds(250, seed=4)
    if not os.path.exists(base + ".done"):
        eb.build_from_collection(col.letters, col.str_off, col.seg_first, base, device=0)
        open(base + ".done", "w").close()
    pats = synth.sample_patterns(col, npat, plen, seed=44)
    del col
    data, off = synth.pack_fixed(pats)
    torch.zeros(1, device="cuda")
    with eb.Index(base, device=0, verify=False) as idx:
        p_data = eb.pinned_empty(data.size, np.uint8); p_data[:] = data
        p_words = eb.pinned_empty(npat, np.uint64); p_inval = eb.pinned_empty((npat + 31) // 32, np.uint32)
        idx.pack_patterns(p_data, plen, nthreads=os.cpu_count() or 1, words=p_words, invalid=p_inval)
        p_counts = eb.pinned_empty(npat, np.uint32)
        p_occ = eb.pinned_empty(npat + 1024, eb.OCC_DTYPE)
        for _ in range(3):
            idx.search_into_packed(p_words, p_inval, plen, npat, p_counts, p_occ, locate=True)
        with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
            idx.search_into_packed(p_words, p_inval, plen, npat, p_counts, p_occ, locate=True)
            torch.cuda.synchronize()
        ev = [e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA]
        ev.sort(key=lambda e: e.time_range.start)
        t0 = ev[0].time_range.start
        busy = {}
        print("# start_us  dur_us  stream  name")
        for e in ev:
            d = e.time_range.end - e.time_range.start
            stream = getattr(e, "stream", None)
            if stream is None:
                stream = -1
            busy[stream] = busy.get(stream, 0) + d
            if d >= 20:
                print("%9.1f %8.1f  %3s  %s" % (e.time_range.start - t0, d, stream, e.name[:70]))
        # one steady-state chunk cycle in full: every device activity and every CUDA runtime call of the host thread
        lo_w, hi_w = t0 + 2750.0, t0 + 3250.0
        print("# window 2750..3250 us: device activities (D) and CUDA runtime calls on the host (H)")
        allev = sorted(prof.events(), key=lambda e: e.time_range.start)
        for e in allev:
            if e.time_range.end < lo_w or e.time_range.start > hi_w:
                continue
            isdev = e.device_type == torch.autograd.DeviceType.CUDA
            if not isdev and not e.name.startswith("cuda"):
                continue
            print("%s %9.1f %8.1f  %s" % ("D" if isdev else "H", e.time_range.start - t0, e.time_range.end - e.time_range.start, e.name[:60]))
        print("# span_us %.1f  activities %d  busy_us per stream %s" % (ev[-1].time_range.end - t0, len(ev), json.dumps({str(k): round(v, 1) for k, v in busy.items()})))


if __name__ == "__main__":
    main()
