#!/usr/bin/env python3
"""Raw host<->device copy bandwidth of the box (pinned memory, CUDA events): the floor of
every `e2e` figure in bench.py.  usage: python tools/pcie_bw.py [MiB]"""
import sys

import torch

mib = int(sys.argv[1]) if len(sys.argv) > 1 else 512
n = mib << 20
h = torch.empty(n, dtype=torch.uint8).pin_memory()
d = torch.empty(n, dtype=torch.uint8, device="cuda")
s2 = torch.cuda.Stream()


def timed(fn, reps=5):
    best = 1e9
    for _ in range(reps):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        a.record()
        fn()
        b.record()
        torch.cuda.synchronize()
        best = min(best, a.elapsed_time(b))
    return best


t = timed(lambda: d.copy_(h, non_blocking=True))
print("H2D  %.1f GB/s" % (n / t / 1e6))
t = timed(lambda: h.copy_(d, non_blocking=True))
print("D2H  %.1f GB/s" % (n / t / 1e6))
h2 = torch.empty(n, dtype=torch.uint8).pin_memory()
d2 = torch.empty(n, dtype=torch.uint8, device="cuda")


def both():
    ev = torch.cuda.Event()
    ev.record()
    with torch.cuda.stream(s2):
        s2.wait_event(ev)
        h2.copy_(d2, non_blocking=True)
    d.copy_(h, non_blocking=True)
    torch.cuda.current_stream().wait_stream(s2)


t = timed(both)
print("H2D+D2H concurrently  %.1f GB/s each direction" % (n / t / 1e6))
for chunk_mib in (1, 4, 16, 64):
    c = chunk_mib << 20

    def chunks():
        for o in range(0, n, c):
            h[o:o + c].copy_(d[o:o + c], non_blocking=True)
    t = timed(chunks, 3)
    print("D2H in %d MiB chunks  %.1f GB/s" % (chunk_mib, n / t / 1e6))
