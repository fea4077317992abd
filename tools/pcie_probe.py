#!/usr/bin/env python
"""PCIe probe: pinned H2D / D2H bandwidth alone and concurrently (the ceiling of the e2e path)."""
import time
import torch
n = 1_240_000_000 // 8
h = torch.empty(n, dtype=torch.int64).pin_memory()
d = torch.empty(n, dtype=torch.int64, device="cuda")
h2 = torch.empty(n // 2, dtype=torch.int64).pin_memory()
d2 = torch.empty(n // 2, dtype=torch.int64, device="cuda")
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
def t(fn, reps=5):
    fn(); torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(reps): fn()
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / reps
def h2d():
    with torch.cuda.stream(s1): d.copy_(h, non_blocking=True)
def d2h():
    with torch.cuda.stream(s2): h2.copy_(d2, non_blocking=True)
def both():
    h2d(); d2h()
a = t(h2d); b = t(d2h); c = t(both)
print(f"H2D {n*8/a/1e9:.1f} GB/s ({a*1e3:.2f} ms for 1.24 GB); D2H {n*4/b/1e9:.1f} GB/s; concurrent: {c*1e3:.2f} ms for 1.24 GB up + 0.62 GB down")
def chunked(k=10):
    m = n // k
    for i in range(k):
        with torch.cuda.stream(s1): d[i*m:(i+1)*m].copy_(h[i*m:(i+1)*m], non_blocking=True)
print(f"H2D in 10 chunks: {t(chunked)*1e3:.2f} ms")
