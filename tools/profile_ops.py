"""One pass over the hot-path ops (for `ncu --metrics gpu__time_duration.sum` launch lists).
    python tools/profile_ops.py [iou|nms|all]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import orientedobjectdetection_b200 as rg
from orientedobjectdetection_b200 import synth

what = sys.argv[1] if len(sys.argv) > 1 else 'all'
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 2
cu = lambda x: torch.from_numpy(x).cuda()
if what in ('iou', 'all'):
    a, b = synth.c3_iou(0)
    ad, bd = cu(a), cu(b)
    for _ in range(reps):
        out = rg.rbbox_overlaps(ad, bd)
    torch.cuda.synchronize()
    print('iou nz', int((out > 0).sum()))
if what in ('nms', 'all'):
    mb, ms = synth.c2_multiclass(0, dense=True)
    mbd, msd = cu(mb), cu(ms)
    for _ in range(reps):
        d, l = rg.multiclass_nms_rotated(mbd, msd, 0.05, dict(iou_thr=0.1), 2000)
    print('c2 dense kept', d.shape[0])
    boxes, scores, ids = synth.c4_rpn_image(0)
    bd, sd, idd = cu(boxes), cu(scores), cu(ids)
    for _ in range(reps):
        d, k = rg.batched_nms(bd, sd, idd, dict(type='nms_rotated', iou_threshold=0.8))
    print('c4 kept', k.numel())
    n = 30000
    b1, s1 = cu(synth.le90_boxes(n, 9, 1024.0)), cu(synth.unique_scores(n, 9))
    for _ in range(reps):
        d, k = rg.nms_rotated(b1, s1, 0.1)
    print('single 30000 kept', k.numel())
    torch.cuda.synchronize()
