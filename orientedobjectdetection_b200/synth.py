"""Seeded synthetic rotated boxes for the five BASELINE.json configs (C1..C5).

Distributions are the ones fixed in SURVEY.md §8(d): le90 boxes
``(cx, cy, w, h, theta)`` with ``w >= h`` (long-edge convention produced by
reference mmrotate/core/bbox/transforms.py:301-331 ``poly2obb_le90``) and
``theta`` in ``[-pi/2, pi/2)`` (range of ``norm_angle(.,'le90')``, reference
transforms.py:864-865).  numpy only: the same arrays feed the CUDA path, the
CPU oracle and bench.py.
"""
from itertools import product
from math import ceil

import numpy as np


def le90_boxes(n, seed, extent=1024.0, w_lo=8.0, w_hi=256.0, aspect_hi=8.0, scale=1.0):
    rng = np.random.default_rng(seed)
    cx = rng.uniform(0.0, extent, n)
    cy = rng.uniform(0.0, extent, n)
    w = np.exp(rng.uniform(np.log(w_lo), np.log(w_hi), n)) * scale
    h = w / rng.uniform(1.0, aspect_hi, n)
    th = rng.uniform(-np.pi / 2, np.pi / 2, n)
    return np.stack([cx, cy, w, h, th], axis=1).astype(np.float32)


def unique_scores(n, seed, lo=0.05):
    """random permutation of linspace(lo+eps, 1, n): unique, all > score_thr"""
    rng = np.random.default_rng(seed + 7919)
    s = np.linspace(lo + 1e-3, 1.0, n, dtype=np.float64)
    return rng.permutation(s).astype(np.float32)


# ---------------------------------------------------------------- C1 / C3
def c1_iou(seed=0, n=1000, m=1000, extent=1024.0):
    """C1: box_iou_rotated on synthetic 1000x1000 le90 boxes."""
    return le90_boxes(n, seed, extent), le90_boxes(m, seed + 1, extent)


def c3_iou(seed=0, n=8192, m=8192, extent=4096.0):
    """C3: 8192x8192 matrix (MaxIoUAssigner / eval shape)."""
    return le90_boxes(n, seed, extent), le90_boxes(m, seed + 1, extent)


# ---------------------------------------------------------------- C2
def c2_multiclass(seed=0, n=2000, num_classes=15, extent=1024.0, dense=False):
    """C2: (multi_bboxes[n,5], multi_scores[n,C+1]) for multiclass_nms_rotated.

    ``dense=False``: about 3 classes per row exceed score_thr=0.05 (N~6000);
    ``dense=True``: every class passes (N = n*C = 30000).  Scores are made
    unique so the score order has no ties.
    """
    rng = np.random.default_rng(seed + 101)
    boxes = le90_boxes(n, seed, extent)
    c = num_classes
    if dense:
        sc = rng.uniform(0.06, 1.0, (n, c))
    else:
        sc = rng.uniform(0.0, 0.045, (n, c))
        hot = rng.random((n, c)) < (3.0 / c)
        sc = np.where(hot, rng.uniform(0.06, 1.0, (n, c)), sc)
    # de-tie: add a distinct tiny ramp
    ramp = rng.permutation(n * c).reshape(n, c) * 1e-7
    sc = (sc + ramp).astype(np.float32)
    bg = np.zeros((n, 1), np.float32)
    return boxes, np.concatenate([sc, bg], axis=1)


# ---------------------------------------------------------------- C4
def c4_rpn_image(seed, per_level=2000, levels=5, extent=1024.0):
    """C4: one image's RPN proposals: 5 levels x 2000, level l scaled 2**l
    (anchor scale 8 x strides 4..64, reference
    configs/oriented_rcnn/oriented_rcnn_r50_fpn_1x_dota_le90.py:31-33)."""
    boxes, ids = [], []
    for lv in range(levels):
        b = le90_boxes(per_level, seed * 31 + lv, extent, w_lo=8.0, w_hi=64.0, scale=2.0 ** lv)
        boxes.append(b)
        ids.append(np.full(per_level, lv, np.int64))
    boxes = np.concatenate(boxes)
    ids = np.concatenate(ids)
    scores = unique_scores(boxes.shape[0], seed, lo=0.0)
    return boxes, scores, ids


# ---------------------------------------------------------------- C5
def slide_windows(width, height, size=1024, step=824):
    """window origins of the reference's slide_window
    (mmrotate/core/patch/split.py:50-61) for one size/step."""
    xn = 1 if width <= size else ceil((width - size) / step + 1)
    xs = [step * i for i in range(xn)]
    if len(xs) > 1 and xs[-1] + size > width:
        xs[-1] = width - size
    yn = 1 if height <= size else ceil((height - size) / step + 1)
    ys = [step * i for i in range(yn)]
    if len(ys) > 1 and ys[-1] + size > height:
        ys[-1] = height - size
    return np.array(list(product(xs, ys)), dtype=np.int64)


def c5_patch_results(seed=0, width=16384, height=16384, size=1024, step=824,
                     dets_per_patch=500, num_classes=15, dup_jitter=0.5):
    """C5: per-patch, per-class det lists in PATCH coordinates, as
    ``merge_results`` receives them (reference
    mmrotate/core/patch/merge_results.py:69-127).  Objects live in global
    coordinates; every patch that contains an object's centre reports it with
    a small jitter, so objects in the overlap bands appear 2-4 times and the
    merge NMS has real duplicates to remove.  Classes are Zipf-skewed.
    Returns (results, offsets, (width, height)).
    """
    rng = np.random.default_rng(seed + 5005)
    offs = slide_windows(width, height, size, step)
    npatch = offs.shape[0]
    # object density chosen so that a patch sees ~dets_per_patch objects
    nobj = int(dets_per_patch * (width / size) * (height / size))
    cx = rng.uniform(0, width, nobj)
    cy = rng.uniform(0, height, nobj)
    w = np.exp(rng.uniform(np.log(8.0), np.log(128.0), nobj))
    h = w / rng.uniform(1.0, 8.0, nobj)
    th = rng.uniform(-np.pi / 2, np.pi / 2, nobj)
    p = 1.0 / np.arange(1, num_classes + 1)
    cls = rng.choice(num_classes, size=nobj, p=p / p.sum())
    base_score = rng.uniform(0.05, 1.0, nobj)
    results = []
    for pi in range(npatch):
        x0, y0 = offs[pi]
        inside = (cx >= x0) & (cx < x0 + size) & (cy >= y0) & (cy < y0 + size)
        idx = np.nonzero(inside)[0]
        k = idx.size
        d = np.empty((k, 6), np.float32)
        d[:, 0] = cx[idx] - x0 + rng.normal(0, dup_jitter, k)
        d[:, 1] = cy[idx] - y0 + rng.normal(0, dup_jitter, k)
        d[:, 2] = w[idx] * (1 + rng.normal(0, 0.01, k))
        d[:, 3] = h[idx] * (1 + rng.normal(0, 0.01, k))
        d[:, 4] = th[idx] + rng.normal(0, 0.01, k)
        d[:, 5] = np.clip(base_score[idx] + rng.normal(0, 0.02, k), 0.05, 1.0) \
            + rng.uniform(0, 1e-4, k)
        results.append([d[cls[idx] == c].copy() for c in range(num_classes)])
    return results, offs, (width, height)


# ---------------------------------------------------------------- stress
def stress_boxes(seed=0):
    """adversarial set for correctness only (SURVEY.md §8(d) 'stress
    variants'): near-duplicates, clamped 1e-3 sizes, axis-aligned angles,
    huge boxes, class-offset-regime coordinates."""
    rng = np.random.default_rng(seed + 4242)
    base = le90_boxes(64, seed, 256.0)
    dup = base.copy()
    dup[:, :2] += rng.normal(0, 1e-4, (64, 2)).astype(np.float32)
    dup[:, 4] += rng.normal(0, 1e-6, 64).astype(np.float32)
    axis = base.copy()
    axis[:, 4] = rng.choice([0.0, np.pi / 2, -np.pi / 2], 64).astype(np.float32)
    tiny = base.copy()
    tiny[:, 2:4] = 1e-3
    huge = base[:4].copy()
    huge[:, 2] = 1.81e7
    huge[:, 3] = 1e7
    far = base.copy()
    far[:, :2] += 3.0e4
    return dict(base=base, dup=dup, axis=axis, tiny=tiny, huge=huge, far=far)


def eval_dataset(seed=0, num_imgs=40, num_classes=15, gts_per_img=60, extent=1024.0, with_ignore=True):
    """detections + annotations in the layout of the reference's eval_rbbox_map
    (mmrotate/core/evaluation/eval_map.py:136-150): per image, ground truths with labels
    (a few of them in the ignore lists) and per class (m, 6) detections made of jittered
    copies of ground truths (some duplicated: the greedy claim matters), wrong-class copies
    and random false alarms."""
    rng = np.random.default_rng(seed + 9001)
    dets, anns = [], []
    for i in range(num_imgs):
        n = int(rng.integers(0, gts_per_img + 1))
        g = le90_boxes(n, seed * 1000 + i, extent, w_lo=12.0, w_hi=160.0)
        lab = rng.integers(0, num_classes, n)
        ign = rng.random(n) < (0.1 if with_ignore else 0.0)
        ann = dict(bboxes=g[~ign], labels=lab[~ign])
        if with_ignore:
            ann['bboxes_ignore'] = g[ign]
            ann['labels_ignore'] = lab[ign]
        anns.append(ann)
        per = []
        for c in range(num_classes):
            src = g[lab == c]
            rep = np.repeat(src, rng.integers(0, 3, src.shape[0]), axis=0)     # 0-2 detections per gt
            jit = rep + rng.normal(0, 1.0, rep.shape).astype(np.float32) * np.array([2, 2, 2, 1, 0.03], np.float32)
            other = g[lab != c][: int(rng.integers(0, 3))]                        # wrong-class hits
            fa = le90_boxes(int(rng.integers(0, 4)), seed * 77 + i * 31 + c, extent)
            b = np.concatenate([jit, other, fa], axis=0).astype(np.float32)
            sc = rng.uniform(0.05, 1.0, b.shape[0]).astype(np.float32)
            if b.shape[0] > 3:
                sc[1] = sc[0]                                                    # a score tie
            per.append(np.concatenate([b, sc[:, None]], axis=1).astype(np.float32))
        dets.append(per)
    return dets, anns
