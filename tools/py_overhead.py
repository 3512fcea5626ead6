"""Where the host time of the small NMS calls goes: cProfile over 300 calls each (C2 sparse multiclass, C4 image batched_nms)."""
import cProfile, os, pstats, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import orientedobjectdetection_b200 as rg
from orientedobjectdetection_b200 import synth
cu = lambda x: torch.from_numpy(x).cuda()
mb, ms = synth.c2_multiclass(0, dense=False)
mbd, msd = cu(mb), cu(ms)
b1, s1, i1 = [cu(x) for x in synth.c4_rpn_image(0)]
cases = {'c2_sparse': lambda: rg.multiclass_nms_rotated(mbd, msd, 0.05, dict(iou_thr=0.1), 2000),
         'c4_image': lambda: rg.batched_nms(b1, s1, i1, dict(type='nms_rotated', iou_threshold=0.8))}
for name, fn in cases.items():
    for _ in range(20):
        fn()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(300):
        fn()
    torch.cuda.synchronize()
    print('== %s: %.1f us per call (300 calls back to back)' % (name, (time.perf_counter() - t0) / 300 * 1e6))
    pr = cProfile.Profile()
    pr.enable()
    for _ in range(300):
        fn()
    pr.disable()
    st = pstats.Stats(pr)
    st.sort_stats('tottime')
    rows = sorted(st.stats.items(), key=lambda kv: -kv[1][2])[:16]
    for (f, ln, fnname), (cc, nc, tt, ct, _) in rows:
        print('   %7.2f us tot  %7.2f us cum  x%-3d  %s:%d %s' % (tt / 300 * 1e6, ct / 300 * 1e6, nc // 300, os.path.basename(f), ln, fnname))
