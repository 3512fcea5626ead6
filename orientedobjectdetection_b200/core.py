"""mmrotate-level seam (SURVEY.md §8(b) "Seam L2"): the functions the
reference exports as ``mmrotate.core.bbox.{rbbox_overlaps, RBboxOverlaps2D}``,
``mmrotate.core.{multiclass_nms_rotated, aug_multiclass_nms_rotated}`` and
``mmrotate.core.merge_results``, with the same signatures, return conventions
and quirks, running on librotgeo.so.
"""
import numpy as np
import torch

from . import ops


# --------------------------------------------------------------------- IoU
def rbbox_overlaps(bboxes1, bboxes2, mode='iou', is_aligned=False):
    """Same contract as reference
    mmrotate/core/bbox/iou_calculators/rotate_iou2d_calculator.py:53-89:
    mode in {'iou','iof'}; last dim 5 (or empty); aligned needs rows == cols;
    empty input returns an (rows, 1) / (rows, cols) tensor (the reference leaves
    it uninitialised, here it is zero-filled); w, h are clamped to >= 1e-3 on a
    private copy — fused into the pack kernel, the caller's tensors are never
    written.
    """
    assert mode in ['iou', 'iof']
    assert (bboxes1.size(-1) == 5 or bboxes1.size(0) == 0)
    assert (bboxes2.size(-1) == 5 or bboxes2.size(0) == 0)
    rows, cols = bboxes1.size(0), bboxes2.size(0)
    if is_aligned:
        assert rows == cols
    if rows * cols == 0:
        return bboxes1.new_zeros((rows, 1)) if is_aligned else bboxes1.new_zeros((rows, cols))
    return ops._overlaps(bboxes1, bboxes2, mode, is_aligned, 1e-3)


class RBboxOverlaps2D(object):
    """2D overlaps calculator, same call contract as reference
    rotate_iou2d_calculator.py:8-50 (registered there as 'RBboxOverlaps2D' in
    mmdet's IOU_CALCULATORS; ``install.patch()`` registers this class under the
    same name when mmdet is importable)."""

    def __call__(self, bboxes1, bboxes2, mode='iou', is_aligned=False, version='oc'):
        assert bboxes1.size(-1) in [0, 5, 6]
        assert bboxes2.size(-1) in [0, 5, 6]
        if bboxes2.size(-1) == 6:
            bboxes2 = bboxes2[..., :5]
        if bboxes1.size(-1) == 6:
            bboxes1 = bboxes1[..., :5]
        # no .contiguous(): the kernels read rows through their stride
        return rbbox_overlaps(bboxes1, bboxes2, mode, is_aligned)

    def __repr__(self):
        return self.__class__.__name__ + '()'


# --------------------------------------------------------------------- NMS
def _iou_thr(nms):
    """``nms.iou_thr`` for ConfigDict-like objects, ``nms['iou_thr']`` for dicts."""
    if isinstance(nms, dict):
        return nms['iou_thr'] if 'iou_thr' in nms else nms['iou_threshold']
    return nms.iou_thr


def multiclass_nms_rotated(multi_bboxes, multi_scores, score_thr, nms, max_num=-1,
                           score_factors=None, return_inds=False):
    """Class-batched rotated NMS, same contract as reference
    mmrotate/core/post_processing/bbox_nms_rotated.py:6-92.

    multi_bboxes (n, C*5) or (n, 5); multi_scores (n, C+1), last column =
    background; candidates are scores > score_thr (tested BEFORE score_factors
    are applied); returns ``(dets (k,6), labels (k,)[, keep])`` where ``keep``
    indexes the FILTERED candidate array.  Classes are separated exactly like
    the reference does — ``label * (max(cx,cy) + max(w,h) + 1)`` added to cx, cy
    in float32 — but the add happens inside the pack kernel and the labels also
    segment the suppression mask, so only same-class tiles are ever computed.
    """
    num_classes = multi_scores.size(1) - 1
    if multi_bboxes.shape[1] > 5:
        bboxes = multi_bboxes.view(multi_scores.size(0), -1, 5)
    else:
        bboxes = multi_bboxes[:, None].expand(multi_scores.size(0), num_classes, 5)
    scores = multi_scores[:, :-1]

    labels = torch.arange(num_classes, dtype=torch.long, device=scores.device)
    labels = labels.view(1, -1).expand_as(scores)
    bboxes = bboxes.reshape(-1, 5)
    scores = scores.reshape(-1)
    labels = labels.reshape(-1)

    valid_mask = scores > score_thr
    if score_factors is not None:
        score_factors = score_factors.view(-1, 1).expand(multi_scores.size(0), num_classes)
        scores = scores * score_factors.reshape(-1)

    inds = valid_mask.nonzero(as_tuple=False).squeeze(1)
    bboxes, scores, labels = bboxes[inds], scores[inds], labels[inds]

    if bboxes.numel() == 0:
        dets = torch.cat([bboxes, scores[:, None]], -1)
        return (dets, labels, inds) if return_inds else (dets, labels)

    max_coordinate = bboxes[:, :2].max() + bboxes[:, 2:4].max()
    offsets = labels.to(bboxes) * (max_coordinate + 1)
    keep = ops._nms_keep(bboxes, scores, _iou_thr(nms), seg_ids=labels, offsets=offsets)
    keep = keep.to(bboxes.device)

    if max_num > 0:
        keep = keep[:max_num]
    bboxes, scores, labels = bboxes[keep], scores[keep], labels[keep]
    dets = torch.cat([bboxes, scores[:, None]], 1)
    return (dets, labels, keep) if return_inds else (dets, labels)


def aug_multiclass_nms_rotated(merged_bboxes, merged_labels, score_thr, nms, max_num, classes):
    """TTA variant, same contract as reference bbox_nms_rotated.py:95-144:
    per class, boxes with score > score_thr go through NMS; the survivors of all
    classes are concatenated class by class and cut to the ``max_num`` best.
    One segmented launch replaces the reference's python loop over classes.
    """
    thr = _iou_thr(nms)
    sel = (merged_bboxes[:, -1] > score_thr) & (merged_labels >= 0) & (merged_labels < classes)
    idx = sel.nonzero(as_tuple=False).squeeze(1)
    if idx.numel() == 0:
        return (merged_bboxes.new_zeros((0, merged_bboxes.size(-1))),
                merged_bboxes.new_zeros((0,), dtype=torch.long))
    cand, lab = merged_bboxes[idx], merged_labels[idx].long()
    keep = ops._nms_keep(cand[:, :5], cand[:, -1], thr, seg_ids=lab).to(cand.device)
    # the reference concatenates per class (ascending), score-descending inside a class
    keep = keep[torch.sort(lab[keep], stable=True).indices]
    bboxes = torch.cat([cand[keep, :5], cand[keep, -1:]], dim=1)
    labels = lab[keep]
    if bboxes.shape[0] > max_num:
        _, _inds = bboxes[:, -1].sort(descending=True, stable=True)
        _inds = _inds[:max_num]
        bboxes, labels = bboxes[_inds], labels[_inds]
    return bboxes, labels


# --------------------------------------------------------------------- patch merge
def translate_bboxes(bboxes, offset):
    """Same as reference mmrotate/core/patch/merge_results.py:7-29 — shifts IN
    PLACE: (n,5) = horizontal x1,y1,x2,y2,score; (n,6) = cx,cy,w,h,theta,score."""
    if bboxes.shape[1] == 5:
        bboxes[:, :4] = bboxes[:, :4] + np.tile(offset, 2)
    elif bboxes.shape[1] == 6:
        bboxes[:, :2] = bboxes[:, :2] + offset
    else:
        raise TypeError('Require the shape of `bboxes` to be (n, 5) or (n, 6),'
                        f' but get `bboxes` with shape being {bboxes.shape}.')
    return bboxes


def merge_results(results, offsets, img_shape, iou_thr=0.1, device='cpu'):
    """Huge-image patch merge, same contract as reference
    mmrotate/core/patch/merge_results.py:69-127 for detection results
    (list over patches of list over classes of (k,6) / (k,5) arrays): every
    patch's boxes are shifted by its window origin (in place, like the
    reference), concatenated per class and de-duplicated by NMS; returns a list
    over classes of numpy arrays ``(k_c, 6)`` in descending-score order.

    All classes go through ONE segmented launch (one H2D, one D2H) instead of a
    per-class copy + NMS + copy.  ``device`` is accepted for compatibility; the
    work always runs on the current CUDA device.  Mask results (tuples) are out
    of scope (no rotated detector of the reference produces them).
    """
    assert len(results) == offsets.shape[0], \
        'The `results` should has the same length with `offsets`.'
    if isinstance(results[0], tuple):
        raise NotImplementedError('mask results are outside the rotated-geometry hot path')
    num_patches = len(results)
    num_classes = len(results[0])

    per_cls = []
    for cls in range(num_classes):
        dets = [translate_bboxes(results[i][cls], offsets[i]) for i in range(num_patches)]
        per_cls.append(np.concatenate(dets, axis=0))

    merged = [None] * num_classes
    rot = [c for c in range(num_classes) if per_cls[c].size and per_cls[c].shape[1] == 6]
    for c in range(num_classes):
        if per_cls[c].size == 0:
            merged[c] = per_cls[c]
        elif per_cls[c].shape[1] == 5:      # horizontal boxes: library NMS, out of scope
            d = torch.from_numpy(per_cls[c])
            out, _ = ops._nms_horizontal(d[:, :-1], d[:, -1], iou_thr)
            merged[c] = out.cpu().numpy()
    if rot:
        counts = [per_cls[c].shape[0] for c in rot]
        allc = np.concatenate([per_cls[c] for c in rot], axis=0)
        seg = np.repeat(np.arange(len(rot), dtype=np.int64), counts)
        dev = ops._cuda_device()
        d = torch.from_numpy(np.ascontiguousarray(allc)).to(dev)
        s = torch.from_numpy(seg).to(dev)
        keep = ops._nms_keep(d[:, :5], d[:, 5], iou_thr, seg_ids=s)
        keep = keep[torch.sort(s[keep], stable=True).indices]       # class-major, score-desc inside
        kept = d[keep].cpu().numpy()
        kseg = s[keep].cpu().numpy()
        bounds = np.searchsorted(kseg, np.arange(len(rot) + 1))
        for j, c in enumerate(rot):
            merged[c] = kept[bounds[j]:bounds[j + 1]].astype(per_cls[c].dtype, copy=False)
    return merged
