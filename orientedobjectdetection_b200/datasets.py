"""Offline patch merge of the DOTA submission path: the reference's
``DOTADataset.merge_det`` / ``_merge_func`` (mmrotate/datasets/dota.py:216-276, 358-382).

The reference groups the per-patch results by original image, then runs one
``nms_rotated`` per (image, class) in a process pool.  Here ALL (image, class) groups of the
dataset go through one segmented launch sequence (``ops._nms_keep`` with one segment per
group); ``nproc`` is accepted and unused.  The rows the reference builds are float64
(``np.zeros(...) + i`` at dota.py:259); the geometry is computed in fp32 on the GPU and the
returned rows are the caller's float64 rows, untouched.
"""
import re
import warnings
from collections import defaultdict

import numpy as np
import torch

from . import ops


def _extract_xy(img_id):
    """reference dota.py:227-247"""
    match = re.compile(r'__(\d+)___(\d+)').search(img_id)
    if match:
        return int(match.group(1)), int(match.group(2))
    warnings.warn("Can't find coordinates in filename, the coordinates will be set to (0,0) by default.",
                  category=Warning)
    return 0, 0


def _nms_groups(label_dets_list, num_classes, iou_thr):
    """label_dets_list[k]: (n_k, 1 + d) rows (label, dets) of group k, where the reference
    takes the first five columns of ``dets`` as the box and the LAST one as the score
    (dota.py:371-380: ``nms_rotated(cls_dets[:, :5], cls_dets[:, -1], iou_thr)``; d is 6 for
    real detections and 5 in the reference's own smoke input, tests/test_data/test_datasets/
    test_dota.py:13-21, whose last box column doubles as the score).
    Returns per group a list over classes of (m, 6) arrays (box + score), score-descending."""
    sizes = [d.shape[0] for d in label_dets_list]
    out = [[None] * num_classes for _ in label_dets_list]
    if sum(sizes) == 0:
        for k, d in enumerate(label_dets_list):
            for c in range(num_classes):
                out[k][c] = d[:0, 1:]      # the reference appends dets[labels == i] itself when empty
        return out
    allr = np.concatenate(label_dets_list, axis=0)
    group = np.repeat(np.arange(len(sizes), dtype=np.int64), sizes)
    labels = allr[:, 0].astype(np.int64)
    inside = (labels >= 0) & (labels < num_classes)
    seg = group * num_classes + np.where(inside, labels, 0)
    idx = np.nonzero(inside)[0]
    # one GROUP of the segmented NMS per (image, class): rows made contiguous per group on the host
    perm = idx[np.argsort(seg[idx], kind='stable')]
    gsizes = np.bincount(seg[perm], minlength=len(sizes) * num_classes)
    dev = ops._cuda_device()
    d = torch.from_numpy(np.ascontiguousarray(allr[perm, 1:6], dtype=np.float32)).to(dev)
    s = torch.from_numpy(np.ascontiguousarray(allr[perm, -1], dtype=np.float32)).to(dev)
    keep, kcounts = ops._nms_grouped(d, s, iou_thr, group_sizes=gsizes)
    gstart = np.concatenate([[0], np.cumsum(gsizes)])[:-1]
    kept_rows = perm[keep.cpu().numpy() + np.repeat(gstart, kcounts)]       # group-major, score order inside
    bounds = np.concatenate([[0], np.cumsum(kcounts)])
    for k in range(len(sizes)):
        for c in range(num_classes):
            lo, hi = bounds[k * num_classes + c], bounds[k * num_classes + c + 1]
            rows = allr[kept_rows[lo:hi]]
            if gsizes[k * num_classes + c] == 0:
                out[k][c] = allr[:0, 1:]       # dota.py:372: an empty class keeps the input width
            else:
                # mmcv.nms_rotated returns cat(dets[keep, :5], scores[keep]): (m, 6) whatever d is
                out[k][c] = np.concatenate([rows[:, 1:6], rows[:, -1:]], axis=1)
    return out


def _merge_func(info, CLASSES, iou_thr):
    """reference dota.py:358-382: ``(img_id, [rows ...]) -> (img_id, [dets per class])``"""
    img_id, label_dets = info
    label_dets = np.concatenate(label_dets, axis=0)
    return img_id, _nms_groups([label_dets], len(CLASSES), iou_thr)[0]


def merge_det(results, img_ids, CLASSES, nproc=4, iou_thr=0.1):
    """reference ``DOTADataset.merge_det`` (dota.py:216-276) as a function of the dataset's
    ``img_ids`` and ``CLASSES``: patch results ``results[idx][cls]`` (k, 6) are shifted by the
    patch origin parsed from the patch name, grouped by original image and de-duplicated per
    class.  Returns ``(id_list, dets_list)`` like ``zip(*merged_results)``."""
    collector = defaultdict(list)
    for idx, img_id in enumerate(img_ids):
        result = results[idx]
        oriname = img_id.split('__', maxsplit=1)[0]
        x, y = _extract_xy(img_id)
        new_result = []
        for i, dets in enumerate(result):
            bboxes, scores = dets[:, :-1], dets[:, [-1]]
            ori_bboxes = bboxes.copy()
            ori_bboxes[..., :2] = ori_bboxes[..., :2] + np.array([x, y], dtype=np.float32)
            labels = np.zeros((bboxes.shape[0], 1)) + i
            new_result.append(np.concatenate([labels, ori_bboxes, scores], axis=1))
        collector[oriname].append(np.concatenate(new_result, axis=0))
    names = list(collector.keys())
    merged = _nms_groups([np.concatenate(collector[n], axis=0) for n in names], len(CLASSES), iou_thr)
    return tuple(names), tuple(merged)
