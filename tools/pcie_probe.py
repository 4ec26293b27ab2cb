#!/usr/bin/env python
"""Raw PCIe rates of the box next to what bench.py's e2e leg achieves: pinned H2D / D2H copies of the e2e sizes, alone and
concurrently on two streams.  python tools/pcie_probe.py > profiles/<round>_pcie.txt"""
import torch

dev = "cuda:0"
torch.cuda.set_device(0)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)


def timed(fn, reps=20):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


for mb in (0.85, 2.36, 18.3, 21.5, 64.0, 256.0):
    nbytes = int(mb * 1e6) // 4 * 4
    h = torch.empty(nbytes, dtype=torch.uint8).pin_memory()
    d = torch.empty(nbytes, dtype=torch.uint8, device=dev)
    t_h2d = timed(lambda: d.copy_(h, non_blocking=True))
    t_d2h = timed(lambda: h.copy_(d, non_blocking=True))
    print(f"{mb:7.2f} MB  H2D {t_h2d * 1e3:8.1f} us = {nbytes / t_h2d / 1e6:6.1f} GB/s   D2H {t_d2h * 1e3:8.1f} us = "
          f"{nbytes / t_d2h / 1e6:6.1f} GB/s")

# the e2e step's pattern: 21.5 MB H2D on one stream, 4.85 MB D2H on another, together
h1 = torch.empty(21_495_808, dtype=torch.uint8).pin_memory()
d1 = torch.empty(21_495_808, dtype=torch.uint8, device=dev)
h2 = torch.empty(4_849_664, dtype=torch.uint8).pin_memory()
d2 = torch.empty(4_849_664, dtype=torch.uint8, device=dev)
side = torch.cuda.Stream()
ev = torch.cuda.Event()


def both():
    with torch.cuda.stream(side):
        h2.copy_(d2, non_blocking=True)
        ev.record(side)
    d1.copy_(h1, non_blocking=True)
    torch.cuda.current_stream().wait_event(ev)


t = timed(both)
print(f"21.5 MB H2D + 4.85 MB D2H on two streams: {t * 1e3:.1f} us")
t = timed(lambda: (d1.copy_(h1, non_blocking=True), h2.copy_(d2, non_blocking=True)))
print(f"the same on one stream: {t * 1e3:.1f} us")
# a host-synchronised step: copy, sync, copy, sync (what a host simulator imposes)
import time
torch.cuda.synchronize()
t0 = time.perf_counter()
for _ in range(20):
    d1.copy_(h1, non_blocking=True)
    torch.cuda.current_stream().synchronize()
t1 = time.perf_counter()
print(f"21.5 MB H2D + host sync, wall: {(t1 - t0) / 20 * 1e6:.1f} us")
