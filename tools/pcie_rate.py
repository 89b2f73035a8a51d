#!/usr/bin/env python
"""Pinned host <-> device copy rate of the box (the ceiling of bench.py e2e: 59 MB in and 10.6 MB out per step)."""
import torch, time
x = torch.empty(59_000_000, dtype=torch.uint8).pin_memory(); y = torch.empty_like(x, device='cuda')
z = torch.empty(10_600_000, dtype=torch.uint8, device='cuda'); zh = torch.empty(10_600_000, dtype=torch.uint8).pin_memory()
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
for mode in ("h2d", "both"):
    torch.cuda.synchronize(); t = time.perf_counter()
    for _ in range(50):
        with torch.cuda.stream(s1): y.copy_(x, non_blocking=True)
        if mode == "both":
            with torch.cuda.stream(s2): zh.copy_(z, non_blocking=True)
    torch.cuda.synchronize(); dt = (time.perf_counter() - t) / 50
    print(mode, "%.3f ms per 59 MB -> %.1f GB/s" % (dt * 1e3, 59e-3 / dt))
