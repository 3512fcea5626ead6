"""mAP evaluation with the detection/ground-truth matching on the GPU (SURVEY.md §8(f)
rank 3).  ``eval_rbbox_map`` keeps the signature and the return value of the reference's
mmrotate/core/evaluation/eval_map.py:126-246; the per-image ``tpfp_default``
(eval_map.py:12-94) that the reference runs once per (class, image) in a 4-process spawn
pool on the CPU is ONE launch over all (class, image) segments (rotgeo_eval_tpfp).
``average_precision`` restates mmdet's function of that name (mmdet/core/evaluation/
mean_ap.py, external dependency of the reference).  No CPU fallback for the matching.
"""
import numpy as np
import torch

from . import _lib
from .ops import _cuda_device, _on_device, _stream


def average_precision(recalls, precisions, mode='area'):
    """mmdet.core.average_precision: 'area' (VOC10+) or '11points' (VOC07); inputs
    (num_scales, num_dets) or (num_dets,)."""
    no_scale = False
    if recalls.ndim == 1:
        no_scale = True
        recalls = recalls[np.newaxis, :]
        precisions = precisions[np.newaxis, :]
    assert recalls.shape == precisions.shape and recalls.ndim == 2
    num_scales = recalls.shape[0]
    ap = np.zeros(num_scales, dtype=np.float32)
    if mode == 'area':
        zeros = np.zeros((num_scales, 1), dtype=recalls.dtype)
        ones = np.ones((num_scales, 1), dtype=recalls.dtype)
        mrec = np.hstack((zeros, recalls, ones))
        mpre = np.hstack((zeros, precisions, zeros))
        for i in range(mpre.shape[1] - 1, 0, -1):
            mpre[:, i - 1] = np.maximum(mpre[:, i - 1], mpre[:, i])
        for i in range(num_scales):
            ind = np.where(mrec[i, 1:] != mrec[i, :-1])[0]
            ap[i] = np.sum((mrec[i, ind + 1] - mrec[i, ind]) * mpre[i, ind + 1])
    elif mode == '11points':
        for i in range(num_scales):
            for thr in np.arange(0, 1 + 1e-3, 0.1):
                precs = precisions[i, recalls[i, :] >= thr]
                prec = precs.max() if precs.size > 0 else 0
                ap[i] += prec
        ap /= 11
    else:
        raise ValueError('Unrecognized mode, only "area" and "11points" are supported')
    if no_scale:
        ap = ap[0]
    return ap


def _flatten(det_results, annotations, num_classes):
    """segment = class * num_imgs + image (class-major, so a class's detections are one
    contiguous slice); inside a segment the ignored gts follow the others (eval_map.py:34-39)"""
    num_imgs = len(det_results)
    det_rows, det_seg, gt_rows, gt_seg, gt_ign = [], [], [], [], []
    for i in range(num_imgs):
        per = det_results[i]
        cnt = [int(np.asarray(d).shape[0]) for d in per]
        if sum(cnt):
            det_rows.append(np.concatenate([np.asarray(d, dtype=np.float32).reshape(-1, 6) for d in per], axis=0))
            det_seg.append(np.repeat(np.arange(num_classes, dtype=np.int64) * num_imgs + i, cnt))
        ann = annotations[i]
        b = np.asarray(ann['bboxes'], dtype=np.float32).reshape(-1, 5)
        lab = np.asarray(ann['labels']).reshape(-1).astype(np.int64)
        ok = (lab >= 0) & (lab < num_classes)
        gt_rows.append(b[ok])
        gt_seg.append(lab[ok] * num_imgs + i)
        gt_ign.append(np.zeros(int(ok.sum()), dtype=np.uint8))
        if ann.get('labels_ignore', None) is not None:
            bi = np.asarray(ann['bboxes_ignore'], dtype=np.float32).reshape(-1, 5)
            li = np.asarray(ann['labels_ignore']).reshape(-1).astype(np.int64)
            ok = (li >= 0) & (li < num_classes)
            gt_rows.append(bi[ok])
            gt_seg.append(li[ok] * num_imgs + i)
            gt_ign.append(np.ones(int(ok.sum()), dtype=np.uint8))
    nseg = num_classes * num_imgs

    def group(rows, seg, width, extra=None):
        rows = np.concatenate(rows, axis=0) if rows else np.zeros((0, width), np.float32)
        seg = np.concatenate(seg) if seg else np.zeros(0, np.int64)
        # stable: the order inside a segment is the order of appearance (dets as given, gts
        # then ignored gts)
        if extra is not None:
            extra = np.concatenate(extra) if extra else np.zeros(0, np.uint8)
            key = seg * 2 + extra
        else:
            key = seg
        o = np.argsort(key, kind='stable')
        off = np.zeros(nseg + 1, dtype=np.int64)
        np.cumsum(np.bincount(seg, minlength=nseg), out=off[1:])
        return np.ascontiguousarray(rows[o]), off, (None if extra is None else np.ascontiguousarray(extra[o]))

    dets, det_off, _ = group(det_rows, det_seg, 6)
    gts, gt_off, ign = group(gt_rows, gt_seg, 5, gt_ign)
    return dets, det_off, gts, gt_off, ign


def tpfp_segments(dets, det_off, gts, gt_off, gt_ignore, iou_thr=0.5):
    """TP / FP flags (uint8 numpy, one per row of ``dets``) for detections and ground truths
    grouped into segments by the two offset arrays: rotgeo_eval_tpfp."""
    dev = _cuda_device()
    nd, ng, nseg = dets.shape[0], gts.shape[0], det_off.shape[0] - 1
    tp = np.zeros(nd, dtype=np.uint8)
    fp = np.zeros(nd, dtype=np.uint8)
    if nd == 0:
        return tp, fp
    L = _lib.load()
    with _on_device(dev):
        d = torch.from_numpy(dets).to(dev)
        g = torch.from_numpy(gts).to(dev)
        do, go = torch.from_numpy(det_off).to(dev), torch.from_numpy(gt_off).to(dev)
        ig = torch.from_numpy(gt_ignore).to(dev) if gt_ignore is not None and gt_ignore.any() else None
        tpd = torch.empty(nd, dtype=torch.uint8, device=dev)
        fpd = torch.empty(nd, dtype=torch.uint8, device=dev)
        nbytes = L.rotgeo_eval_tpfp_workspace_bytes(nd, ng)
        ws = torch.empty(max(nbytes, 16), dtype=torch.uint8, device=dev)
        _lib.check(L.rotgeo_eval_tpfp(d.data_ptr(), nd, 6, do.data_ptr(), g.data_ptr() if ng else None, ng, 5,
                                      go.data_ptr(), ig.data_ptr() if ig is not None else None, nseg,
                                      float(iou_thr), tpd.data_ptr(), fpd.data_ptr(), None, ws.data_ptr(),
                                      nbytes, _stream(dev)), 'rotgeo_eval_tpfp')
        tp, fp = tpd.cpu().numpy(), fpd.cpu().numpy()
    return tp, fp


def eval_rbbox_map(det_results, annotations, scale_ranges=None, iou_thr=0.5, use_07_metric=True,
                   dataset=None, logger=None, nproc=4):
    """Same contract as reference mmrotate/core/evaluation/eval_map.py:126-246:
    ``det_results[img][cls]`` = (m, 6) arrays, ``annotations[img]`` = dict with bboxes (n, 5),
    labels (n,), optionally bboxes_ignore / labels_ignore; returns ``(mean_ap,
    [dict(num_gts, num_dets, recall, precision, ap) per class])``.  ``scale_ranges`` raises
    NotImplementedError like the reference's tpfp_default does for area ranges
    (eval_map.py:56,70); ``nproc`` is accepted and unused (no process pool); ``dataset`` /
    ``logger`` only drive the printed summary."""
    assert len(det_results) == len(annotations)
    if scale_ranges is not None:
        raise NotImplementedError
    num_imgs = len(det_results)
    num_classes = len(det_results[0])
    dets, det_off, gts, gt_off, ign = _flatten(det_results, annotations, num_classes)
    tp_all, fp_all = tpfp_segments(dets, det_off, gts, gt_off, ign, iou_thr)
    # ground truths that count: the not-ignored ones (eval_map.py:190-193)
    seg_of_gt = np.repeat(np.arange(det_off.shape[0] - 1), np.diff(gt_off))
    live = np.bincount(seg_of_gt[ign == 0] if ign is not None else seg_of_gt, minlength=det_off.shape[0] - 1)
    eval_results = []
    eps = np.finfo(np.float32).eps
    for c in range(num_classes):
        lo, hi = det_off[c * num_imgs], det_off[(c + 1) * num_imgs]
        num_gts = int(live[c * num_imgs:(c + 1) * num_imgs].sum())
        cls_dets = dets[lo:hi]
        num_dets = cls_dets.shape[0]
        sort_inds = np.argsort(-cls_dets[:, -1], kind='stable')
        tp = np.cumsum(tp_all[lo:hi][sort_inds].astype(np.float32))
        fp = np.cumsum(fp_all[lo:hi][sort_inds].astype(np.float32))
        recalls = tp / np.maximum(num_gts, eps)
        precisions = tp / np.maximum(tp + fp, eps)
        ap = average_precision(recalls, precisions, 'area' if not use_07_metric else '11points')
        eval_results.append({'num_gts': num_gts, 'num_dets': num_dets, 'recall': recalls,
                             'precision': precisions, 'ap': ap})
    aps = [r['ap'] for r in eval_results if r['num_gts'] > 0]
    mean_ap = np.array(aps).mean().item() if aps else 0.0
    if logger is not None and logger != 'silent':
        print_map_summary(mean_ap, eval_results, dataset, logger=logger)
    return mean_ap, eval_results


def print_map_summary(mean_ap, results, dataset=None, scale_ranges=None, logger=None):
    """plain-text stand-in for reference eval_map.py:249-309 (terminaltables / mmcv.print_log
    are not dependencies of this package)"""
    names = list(dataset) if isinstance(dataset, (list, tuple)) else [str(i) for i in range(len(results))]
    lines = ['%-20s %8s %8s %8s %8s' % ('class', 'gts', 'dets', 'recall', 'ap')]
    for n, r in zip(names, results):
        rec = float(r['recall'][-1]) if r['recall'].size else 0.0
        lines.append('%-20s %8d %8d %8.3f %8.3f' % (n, r['num_gts'], r['num_dets'], rec, float(r['ap'])))
    lines.append('%-20s %8s %8s %8s %8.3f' % ('mAP', '', '', '', mean_ap))
    text = '\n'.join(lines)
    if hasattr(logger, 'info'):
        logger.info('\n' + text)
    else:
        print(text)
