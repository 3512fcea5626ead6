"""One workload, a few calls, for captures under ncu (filter the kernel with -k).
    python tools/run_once.py [c4batch|c5|dense30k|c3dense] [calls]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch
import orientedobjectdetection_b200 as rg
from orientedobjectdetection_b200 import synth, dist as rdist, ops
what = sys.argv[1] if len(sys.argv) > 1 else 'c4batch'
calls = int(sys.argv[2]) if len(sys.argv) > 2 else 3
cu = lambda x: torch.from_numpy(x).cuda()
if what == 'c4batch':
    imgs = [synth.c4_rpn_image(s) for s in range(64)]
    bat = rdist.Batched(cu(np.concatenate([b for b, _, _ in imgs])), cu(np.concatenate([s for _, s, _ in imgs])),
                        cu(np.concatenate([i for _, _, i in imgs])), [b.shape[0] for b, _, _ in imgs])
    fn = lambda: rdist.nms_images(bat, 0.8, 2000)
elif what == 'c5':
    res, offs, shape = synth.c5_patch_results(0)
    for p, o in zip(res, offs):
        for a in p:
            a[:, :2] += o
    per_cls = [np.concatenate([p[c] for p in res], axis=0) for c in range(15)]
    counts = [a.shape[0] for a in per_cls]
    dd = cu(np.concatenate(per_cls, axis=0))
    fn = lambda: ops._nms_grouped(dd[:, :5], dd[:, 5], 0.1, group_sizes=counts)
elif what == 'dense30k':
    bs, sc = cu(synth.le90_boxes(30000, 9, 1024.0)), cu(synth.unique_scores(30000, 9))
    fn = lambda: ops._nms_keep(bs, sc, 0.1)
else:
    a, b = synth.c3_iou(0, extent=1024.0)
    ad, bd = cu(a), cu(b)
    fn = lambda: rg.rbbox_overlaps(ad, bd)
for _ in range(calls):
    fn()
torch.cuda.synchronize()
print('done', what)
