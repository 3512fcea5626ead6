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


_FUSED_WS_LIMIT = 2 << 30


def _multiclass_nms_fused(multi_bboxes, multi_scores, score_thr, iou_thr, max_num, score_factors,
                          return_inds):
    """rotgeo_multiclass_nms; returns None when the shape is outside what it takes."""
    from . import _lib
    n, num_classes = multi_scores.size(0), multi_scores.size(1) - 1
    if n == 0 or num_classes < 1 or num_classes > 64 or n * num_classes >= (1 << 26):
        return None
    if not float(iou_thr) > 0.0:
        return None     # degenerate threshold: answered by the seam (ops._top1), not by the kernels
    if multi_bboxes.shape[1] not in (5, 5 * num_classes):
        return None
    if not multi_scores.is_cuda and not bool((multi_scores[:, :-1] > score_thr).any()):
        return None     # host tensors without a candidate: the empty result needs no device
    L = _lib.load()
    nbytes = L.rotgeo_multiclass_nms_workspace_bytes(n, num_classes)
    if nbytes == 0 or nbytes > _FUSED_WS_LIMIT:
        return None
    dev = ops._cuda_device(multi_bboxes, multi_scores)
    mb = multi_bboxes.detach().to(device=dev, dtype=torch.float32).contiguous()
    ms = multi_scores.detach().to(device=dev, dtype=torch.float32).contiguous()
    sf = None
    if score_factors is not None:
        sf = score_factors.detach().to(device=dev, dtype=torch.float32).reshape(-1).contiguous()
    cap = min(max_num, n * num_classes) if max_num > 0 else n * num_classes
    with ops._on_device(dev):
        stream = ops._stream(dev)
        ws = ops._workspace(dev, stream, 'multiclass', nbytes)      # scratch, reused across calls on this stream
        dets = torch.empty((cap, 6), dtype=torch.float32, device=dev)
        ints = torch.empty(2 * cap + 1, dtype=torch.int64, device=dev)   # labels | flat indices | count: one allocation
        labels, flat, num = ints[:cap], ints[cap:2 * cap], ints[2 * cap:]
        _lib.check(L.rotgeo_multiclass_nms(mb.data_ptr(), mb.shape[1], ms.data_ptr(),
                                           sf.data_ptr() if sf is not None else None, n, num_classes,
                                           float(score_thr), float(iou_thr), cap, dets.data_ptr(),
                                           labels.data_ptr(), flat.data_ptr(), num.data_ptr(),
                                           ws.data_ptr(), ws.numel(), stream),
                   'rotgeo_multiclass_nms')
        k = int(num.item())        # the one device->host read: sizes the outputs
    dets, labels = dets[:k], labels[:k]
    if dets.dtype != multi_bboxes.dtype:
        dets = dets.to(multi_bboxes.dtype)
    if dets.device != multi_bboxes.device:
        dets, labels = dets.to(multi_bboxes.device), labels.to(multi_bboxes.device)
    if not return_inds:
        return dets, labels
    # `keep` of the reference indexes the FILTERED candidate array (row-major order of the
    # candidates): rank of the kept flat index among the valid ones
    valid = (ms[:, :-1] > score_thr).reshape(-1)
    rank = torch.cumsum(valid, 0) - 1
    keep = rank[flat[:k]].to(multi_bboxes.device)
    return dets, labels, keep


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

    The whole body runs as ONE C-ABI call (rotgeo_multiclass_nms: candidate filter,
    sort keys, two radix sorts, device-side plan, pack, mask, scan, emit + output
    gather) and one device->host read of the kept count; the step-by-step torch
    version below it is used only for shapes the fused call does not take
    (more than 64 classes, or a worst-case workspace above 2 GiB).
    """
    fused = _multiclass_nms_fused(multi_bboxes, multi_scores, score_thr, _iou_thr(nms), max_num,
                                  score_factors, return_inds)
    if fused is not None:
        return fused

    num_classes = multi_scores.size(1) - 1
    n = multi_scores.size(0)
    scores2d = multi_scores[:, :-1]
    # candidates: raw score above the threshold (score_factors are applied afterwards, like
    # the reference does)
    valid_mask = scores2d > score_thr
    total = int(valid_mask.sum())
    flat_valid = valid_mask.reshape(-1)
    if hasattr(torch, 'nonzero_static'):
        inds = torch.nonzero_static(flat_valid, size=total).squeeze(1)
    else:
        inds = flat_valid.nonzero(as_tuple=False).squeeze(1)
    rows = torch.div(inds, num_classes, rounding_mode='floor')
    labels = inds - rows * num_classes
    if multi_bboxes.shape[1] > 5:
        bboxes = multi_bboxes.view(n, -1, 5).reshape(-1, 5)[inds]
    else:
        bboxes = multi_bboxes[rows]          # the (n, C, 5) expansion is never materialised
    scores = scores2d.reshape(-1)[inds]
    if score_factors is not None:
        scores = scores * score_factors.view(-1)[rows]

    if bboxes.numel() == 0:
        dets = torch.cat([bboxes, scores[:, None]], -1)
        return (dets, labels, inds) if return_inds else (dets, labels)

    # class separation: label * (max(cx,cy) + max(w,h) + 1) added to cx, cy (id_offsets)
    keep = ops._nms_keep(bboxes, scores, _iou_thr(nms), seg_ids=labels, id_offsets=True)
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


_STAGE = {}


def _stage_rows(rows):
    """pinned float32 staging rows of the merge upload, reused across calls (grown by halves)"""
    buf = _STAGE.get('rows')
    if buf is None or buf.shape[0] < rows:
        buf = torch.empty((max(int(rows * 1.5), 1 << 16), 6), dtype=torch.float32, pin_memory=True)
        _STAGE['rows'] = buf
    return buf


def _merge_results_fast(results, offsets, num_classes, iou_thr):
    """merge_results for the usual input — every class a writable float32 (k,6) array: the
    patches x classes walk (in-place shift + per-class concatenation, reference
    merge_results.py:98-107) runs in the C helper csrc/hostpack.c straight into pinned staging
    rows; one upload, one rotgeo_nms_grouped call with the classes as groups whose resolve
    kernel writes the kept rows itself, one download.  None = input the helper does not take."""
    from . import _hostpack
    if not float(iou_thr) > 0.0:
        return None
    dev = ops._cuda_device()
    offs = np.ascontiguousarray(np.asarray(offsets).reshape(len(results), 2), dtype=np.float64)
    stage = _stage_rows(1)
    got = _hostpack.gather_shift(results, offs, stage.numpy())
    if got is not None and not got[2]:
        stage = _stage_rows(got[0])
        got = _hostpack.gather_shift(results, offs, stage.numpy())
    if got is None:
        return None
    total, counts, _ = got
    empty = np.zeros((0, 6), dtype=np.float32)
    if total == 0:
        return [empty.copy() for _ in range(num_classes)]
    d = stage[:total].to(dev, non_blocking=True)
    groups = [c for c in counts if c]
    got = ops._nms_grouped(d[:, :5], d[:, 5], iou_thr, group_sizes=groups, with_dets=True)
    if got is None:
        raise RuntimeError('rotgeo: the grouped NMS refused a job without segment ids')
    keep, kcounts, dets = got
    if dets is None:
        starts = np.concatenate([[0], np.cumsum(groups)])[:-1]
        rows = keep + torch.repeat_interleave(torch.from_numpy(starts).to(dev),
                                              torch.tensor(kcounts, device=dev), output_size=int(keep.numel()))
        dets = d[rows]
    # the kept rows come back through pinned memory (3 MB for the C5 scene: 60 us instead of ~250 us pageable)
    host = torch.empty(dets.shape, dtype=torch.float32, pin_memory=True)
    host.copy_(dets, non_blocking=True)
    torch.cuda.current_stream(dev).synchronize()
    kept = host.numpy()
    merged, at, j = [], 0, 0
    for c in range(num_classes):
        if counts[c]:
            merged.append(kept[at:at + kcounts[j]])
            at += kcounts[j]
            j += 1
        else:
            merged.append(empty.copy())
    return merged


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

    fast = _merge_results_fast(results, offsets, num_classes, iou_thr)
    if fast is not None:
        return fast
    # arrays the C helper does not take (horizontal (k,5) classes, float64 rows, read-only arrays): numpy walk
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
        # every class is one GROUP of the segmented NMS: one upload, one launch sequence, one
        # download; the kept rows come back class after class, score-descending inside a class
        counts = [per_cls[c].shape[0] for c in rot]
        allc = np.concatenate([per_cls[c] for c in rot], axis=0)
        dev = ops._cuda_device()
        d = torch.from_numpy(np.ascontiguousarray(allc)).to(dev)
        keep, kcounts = ops._nms_grouped(d[:, :5], d[:, 5], iou_thr, group_sizes=counts)
        starts = np.concatenate([[0], np.cumsum(counts)])[:-1]
        rows = keep + torch.repeat_interleave(torch.from_numpy(starts).to(dev),
                                              torch.tensor(kcounts, device=dev), output_size=int(keep.numel()))
        kept = d[rows].cpu().numpy()
        bounds = np.concatenate([[0], np.cumsum(kcounts)])
        for j, c in enumerate(rot):
            merged[c] = kept[bounds[j]:bounds[j + 1]].astype(per_cls[c].dtype, copy=False)
    return merged
