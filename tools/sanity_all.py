"""Every kernel of librotgeo.so once, at small sizes (for compute-sanitizer runs):
    python tools/sanity_all.py"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch
import orientedobjectdetection_b200 as rg
from orientedobjectdetection_b200 import synth, evaluation, losses, datasets, dist as rdist
from oracle import ref, ref_api
cu = lambda x: torch.from_numpy(x).cuda()

a, b = synth.c1_iou(0, 700, 500, 300.0)
m = rg.rbbox_overlaps(cu(a), cu(b)).cpu().numpy()
assert np.abs(m - ref.box_iou_rotated(a, b)).max() <= 1e-5
assert np.abs(rg.box_iou_rotated(cu(a[:500]), cu(b), 'iof', aligned=True).cpu().numpy()
              - ref.box_iou_rotated(a[:500], b, 'iof', True)).max() <= 1e-5
rm, ra, cm, ca = rg.rbbox_overlaps_max(cu(a), cu(b))
assert float((rm.cpu() - torch.from_numpy(m).max(1).values).abs().max()) <= 1e-6
r = rg.MaxIoUAssigner(0.5, 0.4).assign(cu(b), cu(a[:30]))
print('iou/assign ok', int((r.gt_inds > 0).sum()))
for n, thr in ((1500, 0.1), (5200, 0.3)):          # one-CTA path and grid path of the resolve kernel
    bx, sc = synth.le90_boxes(n, 3, 400.0), synth.unique_scores(n, 3)
    _, k = rg.nms_rotated(cu(bx), cu(sc), thr)
    assert k.cpu().tolist() == ref.nms_rotated(bx, sc, thr)[1].tolist()
bx, sc, ids = synth.c4_rpn_image(2, per_level=500)
_, k = rg.batched_nms(cu(bx), cu(sc), torch.from_numpy(ids).cuda(), dict(type='nms_rotated', iou_threshold=0.8))
assert k.cpu().tolist() == ref_api.batched_nms_rotated(torch.from_numpy(bx), torch.from_numpy(sc), torch.from_numpy(ids), 0.8)[1].tolist()
mb, ms = synth.c2_multiclass(1, n=600, num_classes=7, extent=300.0)
d, l = rg.multiclass_nms_rotated(cu(mb), cu(ms), 0.05, dict(iou_thr=0.1), 200)
wd, wl, _ = ref_api.multiclass_nms_rotated(torch.from_numpy(mb), torch.from_numpy(ms), 0.05, 0.1, 200)
assert torch.equal(d.cpu(), wd) and torch.equal(l.cpu(), wl)
print('nms ok', d.shape[0])
dets, anns = synth.eval_dataset(3, num_imgs=8, num_classes=5, gts_per_img=30)
ap, _ = evaluation.eval_rbbox_map(dets, anns)
assert abs(ap - ref_api.eval_rbbox_map(dets, anns)[0]) < 1e-6
pa = cu(a[:300]).requires_grad_(True)
iou = losses.diff_iou_rotated_2d(pa[None], cu(a[:300] + np.float32(3.0))[None])
iou.sum().backward()
assert bool(torch.isfinite(pa.grad).all())
res, offs, shape = synth.c5_patch_results(0, width=2200, height=2200, dets_per_patch=80, num_classes=4)
import copy
mg = rg.merge_results(copy.deepcopy(res), offs, shape, 0.1)
wg = ref_api.merge_results(copy.deepcopy(res), offs, 0.1)
assert all(np.array_equal(x, y) for x, y in zip(mg, wg))
print('eval/diff/merge ok', ap)
