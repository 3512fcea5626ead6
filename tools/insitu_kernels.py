"""In-situ kernel durations (warm caches, back-to-back launches) of the NMS calls through
torch.profiler / CUPTI — ncu's per-kernel times are cold-cache and serialised.
    python tools/insitu_kernels.py"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from torch.profiler import profile, ProfilerActivity
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
