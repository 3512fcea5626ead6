import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch
import orientedobjectdetection_b200 as rg
from oracle import ref
from tests.test_gpu_fuzz import _scene, SIZES
cu = lambda x: torch.from_numpy(np.ascontiguousarray(x)).cuda()
np.set_printoptions(precision=6, suppress=True, linewidth=200)
for seed in range(8):
    rng = np.random.default_rng(3000 + seed)
    n, m = int(rng.choice(SIZES)), int(rng.choice(SIZES))
    a, _ = _scene(rng, n, float(rng.choice([30.0, 400.0])))
    b, _ = _scene(rng, m, float(rng.choice([30.0, 400.0])))
    for mode in ('iou', 'iof'):
        got = rg.box_iou_rotated(cu(a), cu(b), mode=mode).cpu().numpy()
        want = ref.box_iou_rotated(a, b, mode)
        d = np.abs(got - want)
        bad = np.argwhere(d > 1e-5)
        if len(bad):
            i, j = bad[0]
            print('IOU seed', seed, mode, 'bad', len(bad), 'first', i, j, 'got', got[i, j], 'want', want[i, j], 'a', a[i], 'b', b[j])
from oracle import ref_api
for seed in (3, 8, 9, 10, 11):
    rng = np.random.default_rng(1000 + seed)
    n = SIZES[seed % len(SIZES)]
    extent = float(rng.choice([20.0, 100.0, 500.0]))
    thr = float(rng.choice([0.05, 0.1, 0.3, 0.5, 0.8]))
    boxes, scores = _scene(rng, n, extent)
    k = int(rng.choice([1, 3, 17]))
    ids = rng.integers(0, k, size=n).astype(np.int64) * int(rng.choice([1, 5]))
    dets, keep = rg.batched_nms(cu(boxes), cu(scores), cu(ids), dict(type='nms_rotated', iou_threshold=thr))
    tb, ts, ti = torch.from_numpy(boxes), torch.from_numpy(scores), torch.from_numpy(ids)
    wdets, wkeep = ref_api.batched_nms_rotated(tb, ts, ti, thr)
    off = ti.to(tb) * (tb[:, :2].max() + tb[:, 2:4].max() + 1)
    shifted = torch.cat([tb[:, :2] + off[:, None], tb[:, 2:5]], dim=-1).numpy()
    got = keep.cpu().numpy().tolist(); want = wkeep.numpy().tolist()
    print('seed', seed, 'n', n, 'thr', thr, 'k', k, 'ids', np.unique(ids), 'equal', got == want, len(got), len(want))
    if got != want:
        kk = next(i for i in range(min(len(got), len(want))) if got[i] != want[i])
        extra = got[kk] if got[kk] not in want else want[kk]
        prev = np.array(want[:kk])
        io = ref.box_iou_rotated(shifted[prev], shifted[extra:extra + 1])[:, 0]
        ig = rg.box_iou_rotated(cu(shifted[prev]), cu(shifted[extra:extra + 1])).cpu().numpy()[:, 0]
        j = int(np.argmax(np.abs(io - ig)))
        j2 = int(np.argmax(io))
        print('   pos', kk, 'box', extra, 'in got' if extra in got else 'in want', 'id', ids[extra], 'score', scores[extra],
              'max oracle iou', io.max(), 'by', prev[j2], 'id', ids[prev[j2]], 'score', scores[prev[j2]], 'product iou there', ig[j2])
        print('   boxes', boxes[prev[j2]], boxes[extra], 'shifted', shifted[prev[j2]], shifted[extra])

from orientedobjectdetection_b200 import dist as rdist
for seed in (2, 8, 11):
    rng = np.random.default_rng(2000 + seed)
    thr = float(rng.choice([0.1, 0.5, 0.8]))
    max_num = int(rng.choice([5, 50, 100000]))
    images = []
    for _ in range(int(rng.integers(1, 7))):
        n = int(rng.choice([0, 1, 64, 129, 300, 500]))
        boxes, scores = _scene(rng, n, float(rng.choice([50.0, 300.0])))
        ids = rng.integers(0, int(rng.choice([1, 2, 5])), size=n).astype(np.int64)
        images.append((boxes, scores, ids))
    got = rdist.nms_images([tuple(cu(x) for x in im) for im in images], thr, max_num)
    print('IMAGES seed', seed, 'thr', thr, 'max_num', max_num, 'sizes', [im[0].shape[0] for im in images])
    for gi, ((boxes, scores, ids), keep) in enumerate(zip(images, got)):
        if boxes.shape[0] == 0: continue
        tb, ts, ti = torch.from_numpy(boxes), torch.from_numpy(scores), torch.from_numpy(ids)
        wdets, wkeep = ref_api.batched_nms_rotated(tb, ts, ti, thr)
        wkeep = wkeep[:max_num]
        g, w = keep.cpu().numpy().tolist(), wkeep.numpy().tolist()
        single = rg.batched_nms(cu(boxes), cu(scores), cu(ids), dict(type='nms_rotated', iou_threshold=thr))[1].cpu().numpy().tolist()[:max_num]
        if g != w:
            kk = next((i for i in range(min(len(g), len(w))) if g[i] != w[i]), min(len(g), len(w)))
            miss = [x for x in w if x not in g][:5]; extra = [x for x in g if x not in w][:5]
            print('  image', gi, 'n', boxes.shape[0], 'ids', np.unique(ids), 'lens', len(g), len(w), 'pos', kk, 'missing', miss, 'extra', extra, 'single call equal to oracle', single == w)
            for x in miss + extra:
                print('     box', x, boxes[x], 'score', scores[x], 'id', ids[x])
