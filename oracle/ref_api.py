"""ORACLE — TEST INFRASTRUCTURE ONLY.  CPU restatement of the reference's
python front-ends around the native ops, calling the C oracle (oracle/ref.py)
exactly where the reference calls ``mmcv.ops``.  Every function follows the
reference file:line it cites, class separation by the coordinate-offset trick
included, so that keep indices can be compared one-to-one with the CUDA path.
"""
import numpy as np
import torch

from . import ref


def _nms(dets, scores, thr):
    """oracle stand-in for mmcv.ops.nms_rotated on CPU torch tensors"""
    if dets.shape[0] == 0:
        return dets, None
    out, keep = ref.nms_rotated(dets.detach().cpu().numpy(), scores.detach().cpu().numpy(), thr)
    return torch.from_numpy(out).to(dets.dtype), torch.from_numpy(keep)


def rbbox_overlaps(b1, b2, mode='iou', is_aligned=False):
    """reference rotate_iou2d_calculator.py:53-89 (clone, clamp 1e-3, native op)"""
    b1 = b1.detach().clone().float()
    b2 = b2.detach().clone().float()
    b1[:, 2:4].clamp_(min=1e-3)
    b2[:, 2:4].clamp_(min=1e-3)
    out = ref.box_iou_rotated(b1.numpy(), b2.numpy(), mode, is_aligned)
    return torch.from_numpy(out)


def multiclass_nms_rotated(multi_bboxes, multi_scores, score_thr, iou_thr, max_num=-1,
                           score_factors=None):
    """reference bbox_nms_rotated.py:6-92; returns (dets, labels, keep)"""
    num_classes = multi_scores.size(1) - 1
    if multi_bboxes.shape[1] > 5:
        bboxes = multi_bboxes.view(multi_scores.size(0), -1, 5)
    else:
        bboxes = multi_bboxes[:, None].expand(multi_scores.size(0), num_classes, 5)
    scores = multi_scores[:, :-1]
    labels = torch.arange(num_classes, dtype=torch.long).view(1, -1).expand_as(scores)
    bboxes, scores, labels = bboxes.reshape(-1, 5), scores.reshape(-1), labels.reshape(-1)
    valid = scores > score_thr
    if score_factors is not None:
        sf = score_factors.view(-1, 1).expand(multi_scores.size(0), num_classes).reshape(-1)
        scores = scores * sf
    inds = valid.nonzero(as_tuple=False).squeeze(1)
    bboxes, scores, labels = bboxes[inds], scores[inds], labels[inds]
    if bboxes.numel() == 0:
        return torch.cat([bboxes, scores[:, None]], -1), labels, inds
    max_coordinate = bboxes[:, :2].max() + bboxes[:, 2:4].max()
    offsets = labels.to(bboxes) * (max_coordinate + 1)
    for_nms = bboxes.clone()
    for_nms[:, :2] = for_nms[:, :2] + offsets[:, None]
    _, keep = _nms(for_nms, scores, iou_thr)
    if max_num > 0:
        keep = keep[:max_num]
    return torch.cat([bboxes[keep], scores[keep][:, None]], 1), labels[keep], keep


def aug_multiclass_nms_rotated(merged_bboxes, merged_labels, score_thr, iou_thr, max_num, classes):
    """reference bbox_nms_rotated.py:95-144"""
    bboxes, labels = [], []
    for cls in range(classes):
        cb = merged_bboxes[merged_labels == cls]
        inds = cb[:, -1] > score_thr
        if len(inds) == 0:
            continue
        cur = cb[inds, :]
        dets, _ = _nms(cur[:, :5], cur[:, -1], iou_thr)
        if dets.size(0) == 0:
            continue
        bboxes.append(dets)
        labels.append(merged_bboxes.new_full((dets.shape[0],), cls, dtype=torch.long))
    if bboxes:
        bboxes, labels = torch.cat(bboxes), torch.cat(labels)
        if bboxes.shape[0] > max_num:
            _, i = bboxes[:, -1].sort(descending=True, stable=True)
            bboxes, labels = bboxes[i[:max_num]], labels[i[:max_num]]
    else:
        bboxes = merged_bboxes.new_zeros((0, merged_bboxes.size(-1)))
        labels = merged_bboxes.new_zeros((0,), dtype=torch.long)
    return bboxes, labels


def batched_nms_rotated(boxes, scores, idxs, iou_threshold, class_agnostic=False):
    """mmcv.ops.batched_nms with type='nms_rotated' (SURVEY.md §8(a) a9): centre
    offsets idxs*(max+1), one NMS call, score-descending keep."""
    if class_agnostic:
        for_nms = boxes
    else:
        max_coordinate = boxes[..., :2].max() + boxes[..., 2:4].max()
        offsets = idxs.to(boxes) * (max_coordinate + torch.tensor(1).to(boxes))
        for_nms = torch.cat([boxes[..., :2] + offsets[:, None], boxes[..., 2:5]], dim=-1)
    _, keep = _nms(for_nms, scores, iou_threshold)
    return torch.cat([boxes[keep], scores[keep][:, None]], -1), keep


def merge_results(results, offsets, iou_thr=0.1):
    """reference merge_results.py:69-127 (rotated, no masks); mutates results in
    place like the reference's translate_bboxes"""
    num_patches, num_classes = len(results), len(results[0])
    merged = []
    for cls in range(num_classes):
        dets = []
        for i in range(num_patches):
            d = results[i][cls]
            d[:, :2] = d[:, :2] + offsets[i]
            dets.append(d)
        dets = np.concatenate(dets, axis=0)
        if dets.size == 0:
            merged.append(dets)
            continue
        t = torch.from_numpy(dets)
        out, _ = _nms(t[:, :-1], t[:, -1], iou_thr)
        merged.append(out.numpy())
    return merged


def nms_keep_segmented(boxes, scores, iou_threshold, seg_ids=None, offsets=None):
    """oracle stand-in for the product's ``ops._nms_keep`` (same signature):
    greedy NMS inside each segment on boxes whose centres carry ``offsets``,
    kept indices returned in global score-descending order (ties: lower index)."""
    b = boxes.detach().cpu().float().clone()
    s = scores.detach().cpu().float().reshape(-1)
    if offsets is not None:
        b[:, :2] = b[:, :2] + offsets.detach().cpu().float()[:, None]
    seg = torch.zeros(b.shape[0], dtype=torch.long) if seg_ids is None else seg_ids.detach().cpu().long()
    kept = []
    for sid in torch.unique(seg).tolist():
        idx = (seg == sid).nonzero(as_tuple=False).squeeze(1)
        _, k = ref.nms_rotated(b[idx].numpy(), s[idx].numpy(), iou_threshold)
        kept.append(idx[torch.from_numpy(k)])
    kept = torch.cat(kept) if kept else torch.zeros(0, dtype=torch.long)
    order = torch.from_numpy(ref.stable_desc_order(s.numpy()))
    rank = torch.empty_like(order)
    rank[order] = torch.arange(order.numel())
    return kept[torch.argsort(rank[kept])]


def max_iou_assign(overlaps, pos_iou_thr, neg_iou_thr, min_pos_iou=0., gt_max_assign_all=True,
                   match_low_quality=True):
    """mmdet 2.x ``MaxIoUAssigner.assign_wrt_overlaps`` (mmdet/core/bbox/assigners/
    max_iou_assigner.py — external dependency of the reference, pinned by
    mmrotate/__init__.py; the assigner named by the fork's RetinaNet / RPN / R-CNN configs),
    restated on a given (num_gts, num_bboxes) overlap matrix.  Returns (assigned_gt_inds,
    max_overlaps)."""
    num_gts, num_bboxes = overlaps.size(0), overlaps.size(1)
    assigned = overlaps.new_full((num_bboxes,), -1, dtype=torch.long)
    if num_gts == 0 or num_bboxes == 0:
        max_overlaps = overlaps.new_zeros((num_bboxes,))
        if num_gts == 0:
            assigned[:] = 0
        return assigned, max_overlaps
    max_overlaps, argmax_overlaps = overlaps.max(dim=0)
    gt_max_overlaps, gt_argmax_overlaps = overlaps.max(dim=1)
    if isinstance(neg_iou_thr, float):
        assigned[(max_overlaps >= 0) & (max_overlaps < neg_iou_thr)] = 0
    else:
        assigned[(max_overlaps >= neg_iou_thr[0]) & (max_overlaps < neg_iou_thr[1])] = 0
    pos_inds = max_overlaps >= pos_iou_thr
    assigned[pos_inds] = argmax_overlaps[pos_inds] + 1
    if match_low_quality:
        for i in range(num_gts):
            if gt_max_overlaps[i] >= min_pos_iou:
                if gt_max_assign_all:
                    assigned[overlaps[i, :] == gt_max_overlaps[i]] = i + 1
                else:
                    assigned[gt_argmax_overlaps[i]] = i + 1
    return assigned, max_overlaps


def tpfp_default(det_bboxes, gt_bboxes, gt_bboxes_ignore, iou_thr=0.5):
    """reference mmrotate/core/evaluation/eval_map.py:12-94 without area ranges (which the
    reference itself leaves NotImplemented); ties in the score order: lower index first"""
    det_bboxes = np.array(det_bboxes)
    gt_ignore_inds = np.concatenate((np.zeros(gt_bboxes.shape[0], dtype=bool),
                                     np.ones(gt_bboxes_ignore.shape[0], dtype=bool)))
    gt_bboxes = np.vstack((gt_bboxes, gt_bboxes_ignore))
    num_dets, num_gts = det_bboxes.shape[0], gt_bboxes.shape[0]
    tp = np.zeros(num_dets, dtype=np.float32)
    fp = np.zeros(num_dets, dtype=np.float32)
    if num_gts == 0:
        fp[...] = 1
        return tp, fp
    if num_dets == 0:
        return tp, fp
    ious = ref.box_iou_rotated(det_bboxes[:, :5].astype(np.float32), gt_bboxes.astype(np.float32))
    ious_max, ious_argmax = ious.max(axis=1), ious.argmax(axis=1)
    sort_inds = np.argsort(-det_bboxes[:, -1], kind='stable')
    gt_covered = np.zeros(num_gts, dtype=bool)
    for i in sort_inds:
        if ious_max[i] >= iou_thr:
            matched_gt = ious_argmax[i]
            if not gt_ignore_inds[matched_gt]:
                if not gt_covered[matched_gt]:
                    gt_covered[matched_gt] = True
                    tp[i] = 1
                else:
                    fp[i] = 1
        else:
            fp[i] = 1
    return tp, fp


def eval_rbbox_map(det_results, annotations, iou_thr=0.5, use_07_metric=True):
    """reference eval_map.py:126-246 (scale_ranges=None) with mmdet's average_precision
    restated; returns (mean_ap, per-class dicts, per-class (tp, fp) before sorting)"""
    num_imgs, num_classes = len(det_results), len(det_results[0])
    results, flags = [], []
    eps = np.finfo(np.float32).eps
    for c in range(num_classes):
        cls_dets = [np.asarray(r[c], dtype=np.float32).reshape(-1, 6) for r in det_results]
        tps, fps, num_gts = [], [], 0
        for i, ann in enumerate(annotations):
            g = np.asarray(ann['bboxes'], np.float32).reshape(-1, 5)[np.asarray(ann['labels']) == c]
            if ann.get('labels_ignore', None) is not None:
                gi = np.asarray(ann['bboxes_ignore'], np.float32).reshape(-1, 5)[np.asarray(ann['labels_ignore']) == c]
            else:
                gi = np.zeros((0, 5), np.float32)
            t, f = tpfp_default(cls_dets[i], g, gi, iou_thr)
            tps.append(t)
            fps.append(f)
            num_gts += g.shape[0]
        alld = np.vstack(cls_dets)
        tp_raw, fp_raw = np.hstack(tps), np.hstack(fps)
        flags.append((tp_raw, fp_raw))
        order = np.argsort(-alld[:, -1], kind='stable')
        tp, fp = np.cumsum(tp_raw[order]), np.cumsum(fp_raw[order])
        rec = tp / np.maximum(num_gts, eps)
        prec = tp / np.maximum(tp + fp, eps)
        if use_07_metric:
            ap = 0.0
            for thr in np.arange(0, 1 + 1e-3, 0.1):
                p = prec[rec >= thr]
                ap += p.max() if p.size else 0.0
            ap /= 11
        else:
            mrec = np.concatenate([[0.], rec, [1.]])
            mpre = np.concatenate([[0.], prec, [0.]])
            for k in range(mpre.size - 1, 0, -1):
                mpre[k - 1] = max(mpre[k - 1], mpre[k])
            ind = np.where(mrec[1:] != mrec[:-1])[0]
            ap = float(np.sum((mrec[ind + 1] - mrec[ind]) * mpre[ind + 1]))
        results.append({'num_gts': num_gts, 'num_dets': alld.shape[0], 'recall': rec, 'precision': prec, 'ap': ap})
    aps = [r['ap'] for r in results if r['num_gts'] > 0]
    return (float(np.mean(aps)) if aps else 0.0), results, flags


def merge_det(results, img_ids, num_classes, iou_thr=0.1):
    """reference mmrotate/datasets/dota.py:216-276 + _merge_func :358-382 with the oracle NMS:
    float64 rows (label, box, score) exactly as the reference builds them, nms_rotated in the
    dtype it is handed (float64)."""
    import re
    from collections import defaultdict
    collector = defaultdict(list)
    for idx, img_id in enumerate(img_ids):
        oriname = img_id.split('__', maxsplit=1)[0]
        m = re.compile(r'__(\d+)___(\d+)').search(img_id)
        x, y = (int(m.group(1)), int(m.group(2))) if m else (0, 0)
        new_result = []
        for i, dets in enumerate(results[idx]):
            bboxes, scores = dets[:, :-1], dets[:, [-1]]
            ori = bboxes.copy()
            ori[..., :2] = ori[..., :2] + np.array([x, y], dtype=np.float32)
            new_result.append(np.concatenate([np.zeros((bboxes.shape[0], 1)) + i, ori, scores], axis=1))
        collector[oriname].append(np.concatenate(new_result, axis=0))
    ids, out = [], []
    for img_id, label_dets in collector.items():
        label_dets = np.concatenate(label_dets, axis=0)
        labels, dets = label_dets[:, 0], label_dets[:, 1:]
        per = []
        for i in range(num_classes):
            d = dets[labels == i]
            if len(d) == 0:
                per.append(d)
            else:
                nms_dets, _ = ref.nms_rotated(d[:, :5], d[:, -1], iou_thr)
                per.append(nms_dets)
        ids.append(img_id)
        out.append(per)
    return tuple(ids), tuple(out)
