"""Measures pinned H2D / D2H bandwidth on the box (context for bench.py's e2e)."""
import time
import torch

n = 8192 * 8192
dev = torch.device('cuda', 0)
d = torch.empty(n, dtype=torch.float32, device=dev)
h = torch.empty(n, dtype=torch.float32, pin_memory=True)
for name, fn in (('d2h', lambda: h.copy_(d, non_blocking=True)), ('h2d', lambda: d.copy_(h, non_blocking=True))):
    fn(); torch.cuda.synchronize()
    ts = []
    for _ in range(5):
        t0 = time.perf_counter(); fn(); torch.cuda.synchronize(); ts.append(time.perf_counter() - t0)
    print(name, 'GB/s %.1f' % (n * 4 / min(ts) / 1e9), 'ms %.2f' % (1e3 * min(ts)))
t0 = time.perf_counter(); h2 = torch.empty(n, dtype=torch.float32, pin_memory=True); print('pinned alloc 268MB ms %.1f' % (1e3 * (time.perf_counter() - t0)))
del h2
t0 = time.perf_counter(); h2 = torch.empty(n, dtype=torch.float32, pin_memory=True); print('pinned re-alloc ms %.3f' % (1e3 * (time.perf_counter() - t0)))
# two copies in flight on two streams (does the link go faster with 2 DMA engines?)
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
half = n // 2
torch.cuda.synchronize()
t0 = time.perf_counter()
with torch.cuda.stream(s1):
    h[:half].copy_(d[:half], non_blocking=True)
with torch.cuda.stream(s2):
    h[half:].copy_(d[half:], non_blocking=True)
torch.cuda.synchronize()
print('d2h 2 streams GB/s %.1f' % (n * 4 / (time.perf_counter() - t0) / 1e9))
import os
print('cpus', len(os.sched_getaffinity(0)))
