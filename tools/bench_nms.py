"""Rotated-NMS workloads of BASELINE.json at full size (C2, C4, C5) on one GPU:
wall time of the public call with device-resident inputs (C2, C4) or host numpy
lists (C5, as huge_image_demo.py feeds merge_results).
    python tools/bench_nms.py [c2|c4|c5|all] [reps]"""
import os, sys, time, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch
import orientedobjectdetection_b200 as rg
from orientedobjectdetection_b200 import synth, dist as rdist

what = sys.argv[1] if len(sys.argv) > 1 else 'all'
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 5
cu = lambda x: torch.from_numpy(x).cuda()
out = {}


def best(fn, reps=reps):
    fn()
    ts = []
    for _ in range(reps):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        r = fn()
        torch.cuda.synchronize()
        ts.append(time.perf_counter() - t0)
    return min(ts), r


if what in ('c2', 'all'):
    for dense in (False, True):
        mb, ms = synth.c2_multiclass(0, dense=dense)
        mbd, msd = cu(mb), cu(ms)
        n = int((ms[:, :-1] > 0.05).sum())
        t, r = best(lambda: rg.multiclass_nms_rotated(mbd, msd, 0.05, dict(iou_thr=0.1), 2000))
        out['c2_dense' if dense else 'c2_sparse'] = dict(boxes=n, ms=1e3 * t, boxes_per_s=n / t, kept=int(r[0].shape[0]))
if what in ('c4', 'all'):
    imgs = [tuple(cu(x) for x in synth.c4_rpn_image(s)) for s in range(64)]
    n = sum(int(b.shape[0]) for b, _, _ in imgs)
    t, r = best(lambda: rdist.nms_images(imgs, 0.8, 2000))
    out['c4_64img_one_launch'] = dict(boxes=n, ms=1e3 * t, boxes_per_s=n / t, kept=sum(int(k.numel()) for k in r))

    def loop():
        return [rg.batched_nms(b, s, i, dict(type='nms_rotated', iou_threshold=0.8))[1][:2000] for b, s, i in imgs]
    t, r = best(loop, 2)
    out['c4_64img_per_image_calls'] = dict(boxes=n, ms=1e3 * t, boxes_per_s=n / t, kept=sum(int(k.numel()) for k in r))
if what in ('c5', 'all'):
    res, offs, shape = synth.c5_patch_results(0)
    n = sum(a.shape[0] for p in res for a in p)
    per_cls = [sum(p[c].shape[0] for p in res) for c in range(len(res[0]))]

    def run():
        # merge_results shifts the caller's arrays in place: give it a fresh copy each time
        cp = [[a.copy() for a in p] for p in res]
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        m = rg.merge_results(cp, offs, shape, 0.1)
        return time.perf_counter() - t0, m
    run()
    ts = []
    for _ in range(reps):
        t, m = run()
        ts.append(t)
    t = min(ts)
    out['c5_merge_results'] = dict(boxes=n, ms=1e3 * t, boxes_per_s=n / t, kept=sum(int(a.shape[0]) for a in m),
                                   largest_class=max(per_cls))
print(json.dumps(out))
