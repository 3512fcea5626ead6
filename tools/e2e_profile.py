import os, sys, time, cProfile, pstats
sys.path.insert(0, os.getcwd())
import torch
import orientedobjectdetection_b200 as rg
from orientedobjectdetection_b200 import synth
a, b = synth.c3_iou(0)
ah, bh = torch.from_numpy(a).pin_memory(), torch.from_numpy(b).pin_memory()
keep = [rg.rbbox_overlaps(ah, bh), rg.rbbox_overlaps(ah, bh)]; del keep
torch.cuda.synchronize()
t0 = time.perf_counter()
for _ in range(10): r = rg.rbbox_overlaps(ah, bh)
torch.cuda.synchronize()
print('per call ms', (time.perf_counter() - t0) / 10 * 1e3, 'GB/s', 268.435456 / ((time.perf_counter() - t0) / 10 * 1e3))
pr = cProfile.Profile(); pr.enable()
for _ in range(10): r = rg.rbbox_overlaps(ah, bh)
pr.disable()
st = pstats.Stats(pr)
for (f, ln, fn), (cc, nc, tt, ct, _) in sorted(st.stats.items(), key=lambda kv: -kv[1][2])[:12]:
    print('  %8.1f us tot %8.1f us cum x%d %s:%d %s' % (tt / 10 * 1e6, ct / 10 * 1e6, nc // 10, os.path.basename(f), ln, fn))
# bare copy for reference
out = torch.empty((8192, 8192), device='cuda'); host = torch.empty((8192, 8192), pin_memory=True)
host.copy_(out); torch.cuda.synchronize()
t0 = time.perf_counter()
for _ in range(10): host.copy_(out, non_blocking=True)
torch.cuda.synchronize()
print('bare D2H ms', (time.perf_counter() - t0) / 10 * 1e3)
