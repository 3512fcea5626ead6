"""Probe: phase times of nms_resolve_kernel (CTA 0) on the C2 calls, from a library built with
-DROTGEO_RESOLVE_TIMING.    python tools/resolve_phases.py tools/_ablate/librotgeo_rt.so"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from orientedobjectdetection_b200 import _lib
_lib.SO_PATH = os.path.abspath(sys.argv[1])
import orientedobjectdetection_b200 as rg
from orientedobjectdetection_b200 import synth
for dense in (False, True):
    mb, ms = synth.c2_multiclass(0, dense=dense)
    mbd, msd = torch.from_numpy(mb).cuda(), torch.from_numpy(ms).cuda()
    print('C2', 'dense' if dense else 'sparse', '(phases: small segments, large, barrier, count, barrier, offsets, emit)', flush=True)
    for _ in range(4):
        rg.multiclass_nms_rotated(mbd, msd, 0.05, dict(iou_thr=0.1), 2000)
        torch.cuda.synchronize()
b1, s1, i1 = [torch.from_numpy(x).cuda() for x in synth.c4_rpn_image(0)]
print('C4 one image', flush=True)
for _ in range(3):
    rg.batched_nms(b1, s1, i1, dict(type='nms_rotated', iou_threshold=0.8)); torch.cuda.synchronize()
bs, sc = torch.from_numpy(synth.le90_boxes(30000, 9, 1024.0)).cuda(), torch.from_numpy(synth.unique_scores(30000, 9)).cuda()
print('single segment of 30000 boxes, extent 1024 (per-chunk lines: chunks 0, 41, 82, ...)', flush=True)
for _ in range(2):
    rg.nms_rotated(bs, sc, 0.1); torch.cuda.synchronize()
