"""Phase clock of sort_cluster_kernel (probe build with -DROTGEO_SORT_TIMING into tools/_ablate/).
    python tools/sort_phases.py"""
import os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from orientedobjectdetection_b200 import build as b, _lib
out = os.path.join(ROOT, 'tools', '_ablate', 'librotgeo_st.so')
if not os.path.exists(out):
    os.makedirs(os.path.dirname(out), exist_ok=True)
    cmd = [b.nvcc_path(), '-gencode', 'arch=compute_100a,code=sm_100a', '-O3', '-lineinfo', '-std=c++17', '-fmad=true',
           '-DROTGEO_SORT_TIMING', '-Xcompiler', '-fPIC,-O2', '-shared', '-ccbin', '/usr/bin/g++', '-o', out]
    subprocess.check_call(cmd + [os.path.join(b.CSRC, f) for f in b.SOURCES])
_lib.SO_PATH = out
import torch
import orientedobjectdetection_b200 as rg
from orientedobjectdetection_b200 import synth, dist as rdist
cu = lambda x: torch.from_numpy(x).cuda()
b1, s1, i1 = [cu(x) for x in synth.c4_rpn_image(0)]
for _ in range(3):
    print('-- C4 one image'); rg.batched_nms(b1, s1, i1, dict(type='nms_rotated', iou_threshold=0.8)); torch.cuda.synchronize()
mb, ms = synth.c2_multiclass(0, dense=True)
mbd, msd = cu(mb), cu(ms)
for _ in range(2):
    print('-- C2 dense'); rg.multiclass_nms_rotated(mbd, msd, 0.05, dict(iou_thr=0.1), 2000); torch.cuda.synchronize()
imgs = [tuple(cu(x) for x in synth.c4_rpn_image(s)) for s in range(64)]
for _ in range(2):
    print('-- C4 64 images'); rdist.nms_images(imgs, 0.8, 2000); torch.cuda.synchronize()
