"""C1 on the GPU: parity against the fp32 and fp64 oracle + device time of the C-ABI call"""
import os, sys, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch
import orientedobjectdetection_b200 as rg
from orientedobjectdetection_b200 import synth
from oracle import ref
a, b = synth.c1_iou(0)
ad, bd = torch.from_numpy(a).cuda(), torch.from_numpy(b).cuda()
g = rg.box_iou_rotated(ad, bd).cpu().numpy()
o32 = ref.box_iou_rotated(a, b)
o64 = ref.box_iou_rotated(a.astype(np.float64), b.astype(np.float64))
for _ in range(5): rg.box_iou_rotated(ad, bd)
ts = []
for _ in range(20):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); rg.box_iou_rotated(ad, bd); e1.record(); e1.synchronize()
    ts.append(e0.elapsed_time(e1))
print(json.dumps({'c1_max_abs_diff_vs_f32_oracle': float(np.abs(g - o32).max()),
                  'c1_max_abs_diff_vs_f64_oracle': float(np.abs(g - o64).max()),
                  'nonzero_pairs': int((o32 > 0).sum()), 'median_ms': float(np.median(ts)),
                  'mpairs_per_s': 1.0 / (np.median(ts) * 1e-3)}))
