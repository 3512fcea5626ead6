"""Pins the CPU oracle (oracle/) against every golden vector the reference's
own tests hold for this path, against closed-form geometry
and against an independent float64 clipper.  CPU only."""
import json
import math
import os

import numpy as np
import pytest

from oracle import ref
from orientedobjectdetection_b200 import synth
from tests import _geom64


def _load(golden_dir, name):
    return json.load(open(os.path.join(golden_dir, name)))


# ---- reference tests/test_utils/test_overlaps.py:7-22 (asserted for real here)
@pytest.mark.parametrize('dtype', [np.float32, np.float64])
def test_overlaps_golden(golden_dir, dtype):
    g = _load(golden_dir, 'overlaps_golden.json')
    pred = np.array(g['predict'], dtype=dtype)
    gt = np.array(g['gt'], dtype=dtype)
    # rbbox_overlaps clamps w,h to >= 1e-3 on a clone (rotate_iou2d_calculator.py:84-87)
    pred[:, 2:4] = np.maximum(pred[:, 2:4], dtype(1e-3))
    gt[:, 2:4] = np.maximum(gt[:, 2:4], dtype(1e-3))
    ious = ref.box_iou_rotated(pred, gt)
    assert ious.shape == (3, 4)
    assert np.allclose(ious, np.array(g['expect_ious']), atol=g['atol'])
    assert np.all(ious >= 0)


# ---- reference tests/test_data/test_datasets/test_dota.py:51-55 (mAP = 0.7272727)
def _voc07_ap(recalls, precisions):
    ap = 0.0
    for thr in np.arange(0, 1 + 1e-3, 0.1):
        p = precisions[recalls >= thr]
        ap += p.max() if p.size else 0.0
    return ap / 11


def test_dota_map_golden(golden_dir):
    g = _load(golden_dir, 'dota_p0004_golden.json')
    dets = np.array(g['dets'], dtype=np.float64)        # (3,5): last column doubles as score
    gts = np.array(g['gt_obbs_oc_f32'], dtype=np.float32)
    assert gts.shape == (4, 5)
    # tpfp_default (eval_map.py:60-93): float32 IoU on CPU, max/argmax, greedy TP/FP
    ious = ref.box_iou_rotated(dets.astype(np.float32), gts)
    imax, iarg = ious.max(1), ious.argmax(1)
    order = np.argsort(-dets[:, -1])
    covered = np.zeros(4, bool)
    tp = np.zeros(3)
    fp = np.zeros(3)
    for i in order:
        if imax[i] >= 0.5 and not covered[iarg[i]]:
            covered[iarg[i]] = True
            tp[i] = 1
        else:
            fp[i] = 1
    tp, fp = np.cumsum(tp[order]), np.cumsum(fp[order])
    rec = tp / 4
    prec = tp / np.maximum(tp + fp, np.finfo(np.float32).eps)
    ap = _voc07_ap(rec, prec)
    np.testing.assert_almost_equal(ap, g['expected_mAP'], decimal=g['decimals'])
    # the three dets are the OBBs of gts 0..2: IoU ~ 1 on the diagonal, 0 elsewhere
    assert np.all(np.diag(ious[:, :3]) > 0.999)
    assert np.all(iarg == np.arange(3))


# ---- reference tests/test_data/test_pipelines/test_rtransforms.py:140-158 (IoF keeps both)
def test_rrandom_crop_iof_golden():
    h, w = 256 - 10, 256 - 30
    window = np.array([[w / 2, h / 2, w, h, 0]], np.float32)
    for box in ([50., 40., 2., 1., 0.], [50., 42., 3., 1., 0.]):
        for off in ((0, 0), (10, 5), (30, 10)):
            b = np.array([box], np.float32) - np.array([off[0], off[1], 0, 0, 0], np.float32)
            iof = ref.box_iou_rotated(b, window, mode='iof')
            assert iof.shape == (1, 1) and iof[0, 0] > 0.7      # iof_thr default 0.7
            assert abs(iof[0, 0] - 1.0) < 1e-5


# ---- closed forms (SURVEY.md §8(c) "independent cross-checks")
CLOSED = [
    ([0, 0, 1, 1, 0], [0, 0, 1, 1, 0], 1.0),
    ([0, 0, 1, 1, 0], [0, 0, 1, 1, math.pi / 4],
     (2 * math.sqrt(2) - 2) / (2 - (2 * math.sqrt(2) - 2))),
    ([0, 0, 2, 2, 0], [1, 0, 2, 2, 0], 2 / 6),
    ([0, 0, 2, 2, 0], [2, 0, 2, 2, 0], 0.0),                      # touching edge
    ([0, 0, 2, 2, 0], [2, 2, 2, 2, 0], 0.0),                      # touching corner
    ([0, 0, 2, 2, 0], [5, 5, 2, 2, 0.3], 0.0),                    # disjoint
    ([10, 10, 4, 4, 0.7], [10, 10, 2, 2, 0.7], 4 / 16),           # nested, same angle
    ([10, 10, 4, 2, 0], [10, 10, 2, 4, math.pi / 2], 1.0),        # (w,h,t) == (h,w,t+pi/2)
    ([3, 4, 6, 2, 0.4], [3, 4, 6, 2, 0.4 + math.pi], 1.0),        # theta + pi
    ([0, 0, 4, 2, 0], [1, 0.5, 4, 2, 0], (3 * 1.5) / (16 - 3 * 1.5)),
]


@pytest.mark.parametrize('dtype,tol', [(np.float32, 2e-6), (np.float64, 1e-12)])
def test_closed_forms(dtype, tol):
    for b1, b2, want in CLOSED:
        got = ref.pair_iou(b1, b2, dtype=dtype)
        assert abs(got - want) < tol, (b1, b2, got, want)


def test_metamorphic_and_f64_clipper():
    a, b = synth.c1_iou(seed=3, n=60, m=60, extent=200.0)
    o32 = ref.box_iou_rotated(a, b)
    o64 = ref.box_iou_rotated(a.astype(np.float64), b.astype(np.float64))
    ind = np.array([[_geom64.iou(x, y) for y in b] for x in a])
    assert (ind > 0).mean() > 0.1                     # the sample really overlaps
    assert np.abs(o64 - ind).max() < 1e-9             # restatement == independent clipper
    assert np.abs(o32 - ind).max() < 2e-6
    # symmetry, common translation / rotation, mirror
    assert np.abs(ref.box_iou_rotated(b, a).T - o32).max() < 2e-6
    sh = np.array([123.0, -45.0, 0, 0, 0], np.float64)
    assert np.abs(ref.box_iou_rotated(a + sh, b + sh) - o64).max() < 1e-9
    am, bm = a.astype(np.float64).copy(), b.astype(np.float64).copy()
    for m in (am, bm):
        m[:, 0] *= -1
        m[:, 4] *= -1
    assert np.abs(ref.box_iou_rotated(am, bm) - o64).max() < 1e-9
    # IoF divides by the FIRST argument's area
    iof = ref.box_iou_rotated(a.astype(np.float64), b.astype(np.float64), mode='iof')
    ind_iof = np.array([[_geom64.iou(x, y, 'iof') for y in b] for x in a])
    assert np.abs(iof - ind_iof).max() < 1e-9


def test_aligned_and_extra_columns():
    a, b = synth.c1_iou(seed=5, n=50, m=50, extent=100.0)
    full = ref.box_iou_rotated(a, b)
    al = ref.box_iou_rotated(a, b, aligned=True)
    assert al.shape == (50,)
    assert np.array_equal(al, np.diag(full))
    a6 = np.concatenate([a, np.full((50, 1), 0.9, np.float32)], 1)   # score column ignored
    assert np.array_equal(ref.box_iou_rotated(a6, b), full)
    assert np.array_equal(ref.box_iou_rotated(a, b, threads=4), full)


def test_degenerate_area_returns_zero():
    a = np.array([[5, 5, 1e-8, 1e-8, 0.1]], np.float32)
    b = np.array([[5, 5, 10, 10, 0.0]], np.float32)
    assert ref.box_iou_rotated(a, b)[0, 0] == 0.0
    assert ref.box_iou_rotated(b, a)[0, 0] == 0.0


# ---- Spec B: greedy NMS, CPU semantics
def _py_greedy(boxes, scores, thr):
    order = np.argsort(-scores.astype(np.float64), kind='stable')
    dead = np.zeros(len(boxes), bool)
    keep = []
    for a_, i in enumerate(order):
        if dead[i]:
            continue
        keep.append(i)
        for j in order[a_ + 1:]:
            if not dead[j] and np.float32(_geom64.iou(boxes[i], boxes[j])) >= np.float32(thr):
                dead[j] = True
    return np.array(keep, np.int64)


def test_nms_matches_python_greedy():
    boxes = synth.le90_boxes(120, 11, extent=150.0)
    scores = synth.unique_scores(120, 11)
    for thr in (0.1, 0.5, 0.8):
        dets, keep = ref.nms_rotated(boxes, scores, thr)
        want = _py_greedy(boxes, scores, thr)
        assert np.array_equal(keep, want)
        assert dets.shape == (len(keep), 6)
        assert np.array_equal(dets[:, :5], boxes[keep])
        assert np.array_equal(dets[:, 5], scores[keep])
        assert np.all(np.diff(scores[keep]) < 0)       # score-descending


def test_nms_edge_cases():
    d, k = ref.nms_rotated(np.zeros((0, 5), np.float32), np.zeros((0,), np.float32), 0.1)
    assert k is None and d.shape == (0, 5)
    one = np.array([[1, 1, 2, 2, 0]], np.float32)
    d, k = ref.nms_rotated(one, np.array([0.3], np.float32), 0.1)
    assert k.tolist() == [0]
    # exact duplicates: IoU == 1 >= thr, only the best survives; thr == IoU suppresses (>=)
    two = np.array([[0, 0, 2, 2, 0], [1, 0, 2, 2, 0]], np.float32)      # IoU = 1/3
    iou = np.float32(ref.pair_iou(two[0], two[1]))
    _, k = ref.nms_rotated(two, np.array([0.9, 0.8], np.float32), float(iou))
    assert k.tolist() == [0]
    _, k = ref.nms_rotated(two, np.array([0.9, 0.8], np.float32), float(np.nextafter(iou, np.float32(1))))
    assert k.tolist() == [0, 1]
    # ties in score: lower index first (our pinned tie rule)
    _, k = ref.nms_rotated(np.array([[0, 0, 2, 2, 0], [50, 0, 2, 2, 0]], np.float32),
                           np.array([0.5, 0.5], np.float32), 0.1)
    assert k.tolist() == [0, 1]
