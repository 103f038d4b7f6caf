#!/usr/bin/env python
"""Pinned host<->device copy rates of the box (what bounds bench.py's e2e line): 128 MiB blocks, CUDA events."""
import torch

n = 16 << 20  # doubles: 128 MiB, the size of x and of y on C3
h1 = torch.empty(n, dtype=torch.float64).pin_memory()
h2 = torch.empty(n, dtype=torch.float64).pin_memory()
d1 = torch.empty(n, dtype=torch.float64, device="cuda")
d2 = torch.empty(n, dtype=torch.float64, device="cuda")
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()


def timed(fn, reps=10):
    fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / reps


gb = n * 8 / 1e9
t = timed(lambda: d1.copy_(h1, non_blocking=True))
print("H2D 128 MiB pinned: %.3f ms  %.1f GB/s" % (t, gb / t * 1e3))
t = timed(lambda: h2.copy_(d2, non_blocking=True))
print("D2H 128 MiB pinned: %.3f ms  %.1f GB/s" % (t, gb / t * 1e3))


def both():
    e = torch.cuda.Event()
    e.record()
    s1.wait_event(e)
    s2.wait_event(e)
    with torch.cuda.stream(s1):
        d1.copy_(h1, non_blocking=True)
    with torch.cuda.stream(s2):
        h2.copy_(d2, non_blocking=True)
    torch.cuda.current_stream().wait_stream(s1)
    torch.cuda.current_stream().wait_stream(s2)


t = timed(both)
print("H2D + D2H at once : %.3f ms  %.1f GB/s each way" % (t, gb / t * 1e3))
