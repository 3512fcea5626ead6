"""CPU reference timings (oracle = restated mmcv CPU algorithm, single thread like mmcv) for the
BASELINE.md table: C1 IoU, C2 multiclass NMS, C4 RPN batched NMS (4 of 64 images, scaled),
C5 patch merge (the 5 smallest + 2 mid classes timed, scaled by sum n_c^2... stated).
    python tools/cpu_nms_baseline.py"""
import json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch
from oracle import ref, ref_api
from orientedobjectdetection_b200 import synth

out = {'cores': len(os.sched_getaffinity(0))}


def best(fn, reps=3):
    fn()
    ts = []
    for _ in range(reps):
        t0 = time.perf_counter()
        fn()
        ts.append(time.perf_counter() - t0)
    return min(ts)


a, b = synth.c1_iou(0)
t1 = best(lambda: ref.box_iou_rotated(a, b, threads=1))
tn = best(lambda: ref.box_iou_rotated(a, b, threads=out['cores']))
out['c1_iou_1000x1000'] = {'mpairs_per_s_1thread': 1.0 / t1, 'mpairs_per_s_all_cores': 1.0 / tn}
for dense in (False, True):
    mb, ms = synth.c2_multiclass(0, dense=dense)
    tb, ts = torch.from_numpy(mb), torch.from_numpy(ms)
    n = int((ms[:, :-1] > 0.05).sum())
    t = best(lambda: ref_api.multiclass_nms_rotated(tb, ts, 0.05, 0.1, 2000), 2)
    out['c2_dense' if dense else 'c2_sparse'] = {'boxes': n, 'ms': 1e3 * t, 'boxes_per_s': n / t}
imgs = [synth.c4_rpn_image(s) for s in range(4)]
def c4():
    for bx, sc, ids in imgs:
        ref_api.batched_nms_rotated(torch.from_numpy(bx), torch.from_numpy(sc), torch.from_numpy(ids), 0.8)
t = best(c4, 2)
out['c4_4_of_64_images'] = {'boxes': 40000, 'ms': 1e3 * t, 'boxes_per_s': 40000 / t,
                            'note': '4 of the 64 images timed; images are independent, scale linearly'}
res, offs, shape = synth.c5_patch_results(0)
ncls = len(res[0])
per = []
for c in range(ncls):
    d = np.concatenate([res[p][c] + np.array([offs[p][0], offs[p][1], 0, 0, 0, 0], np.float32) for p in range(len(res))])
    per.append(d.astype(np.float32))
sizes = [d.shape[0] for d in per]
timed = [c for c in range(ncls) if sizes[c] <= 14000]
t0 = time.perf_counter()
for c in timed:
    ref.nms_rotated(per[c][:, :5], per[c][:, 5], 0.1)
t = time.perf_counter() - t0
nb = sum(sizes[c] for c in timed)
# the greedy CPU loop costs ~ kept x n per class: scale the untimed classes by n^2
scale = sum(s * s for s in sizes) / sum(sizes[c] ** 2 for c in timed)
out['c5_merge'] = {'class_sizes': sizes, 'timed_classes': timed, 'timed_boxes': nb, 'timed_s': t,
                   'extrapolated_full_s': t * scale, 'boxes_per_s_extrapolated': sum(sizes) / (t * scale),
                   'note': 'classes of <= 14000 boxes timed in full, the rest extrapolated by sum n_c^2'}
print(json.dumps(out))
