"""Independent greedy rotated NMS (test helper) — the only pin NMS semantics can get in this
image: the reference holds no NMS vector and mmcv cannot be installed, so
the oracle's ``nms_rotated`` (oracle/rotgeo_ref_impl.inc) is cross-checked against a second
implementation that shares neither code nor algorithm with it:

  * geometry: ``tests/_geom64.py`` (float64 Sutherland-Hodgman + shoelace, pure python) or
    OpenCV's ``cv2.rotatedRectangleIntersection`` + ``cv2.contourArea`` (third-party, fp32
    inside) — the oracle uses mmcv's edge/edge hits + Graham scan + triangle fan;
  * greedy loop: a plain python loop over the score order with the reference's CPU test
    ``IoU >= thr`` (mmcv nms_rotated_cpu, SURVEY.md Spec B), pairs pre-filtered by their
    circumscribed circles (disjoint circles => IoU exactly 0 < thr for thr > 0).

``greedy_nms`` also returns every evaluated pair whose IoU lies within ``band`` of the
threshold: the only pairs at which two correct implementations may legitimately disagree.
"""
import math

import numpy as np

from tests import _geom64


def iou_cv2(b1, b2):
    import cv2
    r1 = ((float(b1[0]), float(b1[1])), (float(b1[2]), float(b1[3])), math.degrees(float(b1[4])))
    r2 = ((float(b2[0]), float(b2[1])), (float(b2[2]), float(b2[3])), math.degrees(float(b2[4])))
    ret, pts = cv2.rotatedRectangleIntersection(r1, r2)
    if ret == 0 or pts is None or len(pts) < 3:
        return 0.0
    inter = float(cv2.contourArea(cv2.convexHull(pts, returnPoints=True)))
    a1, a2 = float(b1[2]) * float(b1[3]), float(b2[2]) * float(b2[3])
    return inter / (a1 + a2 - inter)


def iou_geom64(b1, b2):
    return _geom64.iou(b1, b2)


def greedy_nms(boxes, scores, thr, iou_fn=iou_geom64, band=1e-5):
    """returns (keep indices in descending-score order, [(i, j, iou) inside the band])"""
    boxes = np.asarray(boxes, dtype=np.float64)
    scores = np.asarray(scores, dtype=np.float64)
    n = boxes.shape[0]
    order = np.argsort(-scores, kind='stable')
    radius = 0.5 * np.hypot(boxes[:, 2], boxes[:, 3]) * (1 + 1e-9) + 1e-9
    keep, near = [], []
    kept_xy = np.zeros((n, 2))
    kept_r = np.zeros(n)
    for i in order:
        suppressed = False
        k = len(keep)
        if k:
            d2 = ((kept_xy[:k] - boxes[i, :2]) ** 2).sum(1)
            cand = np.nonzero(d2 <= (kept_r[:k] + radius[i]) ** 2)[0]
            for c in cand:                       # kept boxes in descending score order
                j = keep[c]
                v = iou_fn(boxes[j], boxes[i])
                if abs(v - thr) <= band:
                    near.append((int(j), int(i), float(v)))
                if v >= thr:
                    suppressed = True
                    break
        if not suppressed:
            kept_xy[k] = boxes[i, :2]
            kept_r[k] = radius[i]
            keep.append(int(i))
    return keep, near
