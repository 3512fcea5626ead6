"""In-situ kernel durations (warm caches, back-to-back launches) of the NMS calls through
torch.profiler / CUPTI — ncu's per-kernel times are cold-cache and serialised.
    python tools/insitu_kernels.py"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from torch.profiler import profile, ProfilerActivity
from orientedobjectdetection_b200 import _lib as _l
if os.environ.get('ROTGEO_SO'):
    _l.SO_PATH = os.path.abspath(os.environ['ROTGEO_SO'])
import orientedobjectdetection_b200 as rg
from orientedobjectdetection_b200 import synth, dist as rdist, ops
cu = lambda x: torch.from_numpy(x).cuda()


def run(name, fn, reps=10):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    with profile(activities=[ProfilerActivity.CUDA]) as prof:
        for _ in range(reps):
            fn()
        torch.cuda.synchronize()
    rows = [(e.key, e.device_time_total / reps, e.count / reps) for e in prof.key_averages()]
    rows.sort(key=lambda r: -r[1])
    tot = sum(r[1] for r in rows)
    print('== %s: %.1f us of GPU time per call' % (name, tot))
    for k, t, c in rows[:9]:
        print('   %7.1f us  x%.0f  %s' % (t, c, k[:90]))


for dense in (False, True):
    mb, ms = synth.c2_multiclass(0, dense=dense)
    mbd, msd = cu(mb), cu(ms)
    run('C2 %s' % ('dense' if dense else 'sparse'), lambda: rg.multiclass_nms_rotated(mbd, msd, 0.05, dict(iou_thr=0.1), 2000))
b1, s1, i1 = [cu(x) for x in synth.c4_rpn_image(0)]
run('C4 one image batched_nms', lambda: rg.batched_nms(b1, s1, i1, dict(type='nms_rotated', iou_threshold=0.8)))
imgs = [tuple(cu(x) for x in synth.c4_rpn_image(s)) for s in range(64)]
run('C4 64 images nms_images', lambda: rdist.nms_images(imgs, 0.8, 2000), 5)
bs, sc = cu(synth.le90_boxes(30000, 9, 1024.0)), cu(synth.unique_scores(30000, 9))
run('single segment 30000', lambda: ops._nms_keep(bs, sc, 0.1), 5)
# C5: ~200k detections, one GROUP per class (class-contiguous rows, as merge_results uploads them)
import numpy as np, time
res, offs, shape = synth.c5_patch_results(0)
for p, o in zip(res, offs):
    for a in p:
        a[:, :2] += o
per_cls = [np.concatenate([p[c] for p in res], axis=0) for c in range(15)]
counts = [a.shape[0] for a in per_cls]
dd = cu(np.concatenate(per_cls, axis=0))
run('C5 200k, classes as groups', lambda: ops._nms_grouped(dd[:, :5], dd[:, 5], 0.1, group_sizes=counts), 5)


def wall(name, fn, reps=20):
    fn()
    ts = []
    for _ in range(reps):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        fn()
        torch.cuda.synchronize()
        ts.append(time.perf_counter() - t0)
    print('wall %-40s best %.3f ms  median %.3f ms' % (name, 1e3 * min(ts), 1e3 * sorted(ts)[len(ts) // 2]))


for dense in (False, True):
    mb, ms = synth.c2_multiclass(0, dense=dense)
    mbd, msd = cu(mb), cu(ms)
    wall('C2 %s' % ('dense' if dense else 'sparse'), lambda: rg.multiclass_nms_rotated(mbd, msd, 0.05, dict(iou_thr=0.1), 2000))
cfg = dict(type='nms_rotated', iou_threshold=0.8)
wall('C4 one image', lambda: rg.batched_nms(b1, s1, i1, cfg))
wall('C4 64 images one call', lambda: rdist.nms_images(imgs, 0.8, 2000), 5)
wall('C4 64 per-image calls', lambda: [rg.batched_nms(b, s, i, cfg) for b, s, i in imgs], 3)
wall('C4 8 images one call', lambda: rdist.nms_images(imgs[:8], 0.8, 2000), 10)
wall('single segment 30000', lambda: ops._nms_keep(bs, sc, 0.1), 5)
wall('C5 groups device-resident', lambda: ops._nms_grouped(dd[:, :5], dd[:, 5], 0.1, group_sizes=counts), 5)
