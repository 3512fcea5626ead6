"""Where the wall time of the batched NMS calls goes (C4 64-image batch, C5 class-batched):
the torch preamble of nms_images, the two C-ABI calls of ops._nms_keep (prepare = sorts + the
sync, prepared = pack/mask/resolve), the torch postamble.
    python tools/breakdown_nms.py"""
import os, sys, time, json, ctypes
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch
import orientedobjectdetection_b200 as rg
from orientedobjectdetection_b200 import synth, dist as rdist, ops, _lib

cu = lambda x: torch.from_numpy(x).cuda()
out = {}


def best(fn, reps=7):
    fn()
    ts = []
    for _ in range(reps):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        r = fn()
        torch.cuda.synchronize()
        ts.append(time.perf_counter() - t0)
    return 1e3 * min(ts), r


imgs = [tuple(cu(x) for x in synth.c4_rpn_image(s)) for s in range(64)]
t_all, _ = best(lambda: rdist.nms_images(imgs, 0.8, 2000))
captured = {}


def spy(boxes, scores, thr, seg_ids=None, offsets=None):
    captured['args'] = (boxes, scores, thr, seg_ids, offsets)
    return ops._nms_keep(boxes, scores, thr, seg_ids=seg_ids, offsets=offsets)


rdist.nms_images(imgs, 0.8, 2000, keep_fn=spy)
b, s, thr, seg, off = captured['args']
t_keep, keep = best(lambda: ops._nms_keep(b, s, thr, seg_ids=seg, offsets=off))
out['c4_64img'] = dict(total_ms=t_all, nms_keep_ms=t_keep, torch_pre_post_ms=t_all - t_keep)

# inside _nms_keep: prepare vs prepared
L = _lib.load()
n = b.shape[0]
st = torch.cuda.current_stream().cuda_stream
segc = seg.contiguous()
pbytes = L.rotgeo_nms_prepare_bytes(n)
prep = torch.empty(pbytes, dtype=torch.uint8, device='cuda')
counts = np.empty(min(n, 16384), dtype=np.int64)
nseg = ctypes.c_int64(0)


def do_prepare():
    return L.rotgeo_nms_prepare(b.data_ptr(), n, b.stride(0), s.data_ptr(), segc.data_ptr(), prep.data_ptr(),
                                pbytes, counts.ctypes.data, counts.size, ctypes.byref(nseg), st)


t_prep, rc = best(do_prepare)
assert rc == 0
out['c4_64img']['prepare_ms'] = t_prep
out['c4_64img']['prepared_and_rest_ms'] = t_keep - t_prep
out['c4_64img']['segments'] = int(nseg.value)

# C5 device resident
res, offs, shape = synth.c5_patch_results(0)
ncls = len(res[0])
per = [np.concatenate([res[i][c] + np.array([offs[i][0], offs[i][1], 0, 0, 0, 0], dtype=np.float32)
                       for i in range(len(res))]) for c in range(ncls)]
allc = cu(np.concatenate(per).astype(np.float32))
segs = cu(np.repeat(np.arange(ncls, dtype=np.int64), [p.shape[0] for p in per]))
bb, ss = allc[:, :5].contiguous(), allc[:, 5].contiguous()
t5, k5 = best(lambda: ops._nms_keep(bb, ss, 0.1, seg_ids=segs))
n5 = bb.shape[0]
pb5 = L.rotgeo_nms_prepare_bytes(n5)
prep5 = torch.empty(pb5, dtype=torch.uint8, device='cuda')


def do_prepare5():
    return L.rotgeo_nms_prepare(bb.data_ptr(), n5, bb.stride(0), ss.data_ptr(), segs.data_ptr(), prep5.data_ptr(),
                                pb5, counts.ctypes.data, counts.size, ctypes.byref(nseg), st)


t5p, rc = best(do_prepare5)
out['c5_device'] = dict(boxes=n5, nms_keep_ms=t5, prepare_ms=t5p, prepared_and_rest_ms=t5 - t5p, kept=int(k5.numel()))

# one RPN image through batched_nms
b1, s1, i1 = imgs[0]
t1, _ = best(lambda: rg.batched_nms(b1, s1, i1, dict(type='nms_rotated', iou_threshold=0.8)))
t1k, _ = best(lambda: ops._nms_keep(b1, s1, 0.8, seg_ids=i1, id_offsets=True))
out['c4_one_image'] = dict(batched_nms_ms=t1, nms_keep_ms=t1k)
# single segment 30000
bs, sc = cu(synth.le90_boxes(30000, 9, 1024.0)), cu(synth.unique_scores(30000, 9))
ts, _ = best(lambda: ops._nms_keep(bs, sc, 0.1))
out['single_30000'] = dict(nms_keep_ms=ts)
print(json.dumps(out, indent=1))
