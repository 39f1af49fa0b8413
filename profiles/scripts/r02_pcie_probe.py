"""Round 2: what the box's PCIe link gives pinned copies (the bound of the e2e leg): D2H and H2D alone, both at once, and D2H split over two streams."""
import json
import torch

def timed(fn, n=10):
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    fn(); torch.cuda.synchronize()
    e0.record()
    for _ in range(n):
        fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n

def main():
    out = {}
    for mb in (16, 64, 256):
        nb = mb << 20
        d = torch.empty(nb, dtype=torch.uint8, device="cuda"); d2 = torch.empty(nb, dtype=torch.uint8, device="cuda")
        h = torch.empty(nb, dtype=torch.uint8).pin_memory(); h2 = torch.empty(nb, dtype=torch.uint8).pin_memory()
        s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
        out["d2h_%dMB_GBps" % mb] = round(nb / timed(lambda: h.copy_(d, non_blocking=True)) / 1e6, 1)
        out["h2d_%dMB_GBps" % mb] = round(nb / timed(lambda: d.copy_(h, non_blocking=True)) / 1e6, 1)
        def both():
            with torch.cuda.stream(s1): h.copy_(d, non_blocking=True)
            with torch.cuda.stream(s2): d2.copy_(h2, non_blocking=True)
            torch.cuda.current_stream().wait_stream(s1); torch.cuda.current_stream().wait_stream(s2)
        out["both_%dMB_each_GBps" % mb] = round(nb / timed(both) / 1e6, 1)
        def split():
            half = nb // 2
            with torch.cuda.stream(s1): h[:half].copy_(d[:half], non_blocking=True)
            with torch.cuda.stream(s2): h[half:].copy_(d[half:], non_blocking=True)
            torch.cuda.current_stream().wait_stream(s1); torch.cuda.current_stream().wait_stream(s2)
        out["d2h_split2_%dMB_GBps" % mb] = round(nb / timed(split) / 1e6, 1)
    print(json.dumps(out))

main()
