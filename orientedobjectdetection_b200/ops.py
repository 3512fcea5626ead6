"""Operator-level seam: the three names the reference binds with
``from mmcv.ops import box_iou_rotated, nms_rotated, batched_nms`` (reference
mmrotate/core/bbox/iou_calculators/rotate_iou2d_calculator.py:2,
mmrotate/core/post_processing/bbox_nms_rotated.py:3,
mmrotate/core/patch/merge_results.py:4, mmrotate/datasets/dota.py:16,
mmrotate/models/dense_heads/oriented_rpn_head.py:6, rotated_rpn_head.py:7) with
the same signatures, argument meaning, return conventions and error behaviour
(§8(b)).

Everything computes on the GPU through librotgeo.so.  CPU tensors (three
reference call sites pass them: eval_map.py:60-62, transforms.py:369-371,
merge_results.py:115 with device='cpu') are moved to the current CUDA device
and the result is returned on the caller's device — a transfer, never a CPU
implementation.  No GPU or no library => RuntimeError.
"""
import ctypes

import numpy as np
import torch

from . import _lib

_MODE = {'iou': 0, 'iof': 1}


def _cuda_device(*tensors):
    if torch.cuda._is_in_bad_fork():
        # e.g. a DataLoader / multiprocessing-pool worker forked after the parent touched CUDA
        # (the reference reaches RRandomCrop and dota._merge_func that way)
        raise RuntimeError('rotgeo: called in a forked child of a process that initialised CUDA; '
                           'the GPU cannot be used here and there is no CPU fallback. Use the '
                           "'spawn' start method, or leave this call site on the reference's CPU op")
    for t in tensors:
        if t is not None and t.is_cuda:
            return t.device
    if not torch.cuda.is_available():
        raise RuntimeError('rotgeo: no CUDA device available and there is no CPU fallback')
    return torch.device('cuda', torch.cuda.current_device())


def _rows_f32(t, dev):
    """(n, >=5) tensor -> float32 CUDA tensor whose rows are readable with a
    row stride (only the first five columns are ever read)."""
    t = t.detach()
    if t.dtype != torch.float32 or t.device != dev:
        t = t.to(device=dev, dtype=torch.float32)
    if t.dim() != 2 or t.stride(1) != 1 or t.stride(0) < 5:
        t = t.contiguous()
    return t


def _stream(dev):
    return torch.cuda.current_stream(dev).cuda_stream


def _overlaps(bboxes1, bboxes2, mode, aligned, min_wh):
    """shared body of box_iou_rotated / rbbox_overlaps (min_wh fuses the clamp
    of rotate_iou2d_calculator.py:84-87 into the pack kernel)."""
    assert mode in _MODE
    rows, cols = bboxes1.size(0), bboxes2.size(0)
    if aligned:
        assert rows == cols
    out_shape = (rows,) if aligned else (rows, cols)
    if rows * cols == 0:
        return bboxes1.new_zeros(out_shape)
    dev = _cuda_device(bboxes1, bboxes2)
    b1, b2 = _rows_f32(bboxes1, dev), _rows_f32(bboxes2, dev)
    L = _lib.load()
    with torch.cuda.device(dev):
        out = torch.empty(out_shape, dtype=torch.float32, device=dev)
        if aligned:
            _lib.check(L.rotgeo_iou_aligned(b1.data_ptr(), b2.data_ptr(), rows, b1.stride(0),
                                            b2.stride(0), _MODE[mode], float(min_wh),
                                            out.data_ptr(), _stream(dev)), 'rotgeo_iou_aligned')
        else:
            nbytes = L.rotgeo_iou_workspace_bytes(rows, cols)
            ws = torch.empty(nbytes, dtype=torch.uint8, device=dev)
            _lib.check(L.rotgeo_iou_pairwise(b1.data_ptr(), rows, b1.stride(0), b2.data_ptr(),
                                             cols, b2.stride(0), _MODE[mode], float(min_wh),
                                             out.data_ptr(), cols, ws.data_ptr(), nbytes,
                                             _stream(dev)), 'rotgeo_iou_pairwise')
    if not bboxes1.is_cuda:
        # CPU caller (eval_map.py:60-62, transforms.py:369-371): result goes back
        # through a pinned buffer (torch's caching host allocator reuses it)
        host = torch.empty(out_shape, dtype=torch.float32, pin_memory=True)
        host.copy_(out, non_blocking=True)
        torch.cuda.current_stream(dev).synchronize()
        out = host
    if out.dtype != bboxes1.dtype:
        out = out.to(dtype=bboxes1.dtype)
    return out


def box_iou_rotated(bboxes1, bboxes2, mode='iou', aligned=False, clockwise=True):
    """Drop-in for ``mmcv.ops.box_iou_rotated``.

    ``bboxes1`` (N,5+) and ``bboxes2`` (M,5+) rows ``(cx, cy, w, h, theta)``;
    returns (N, M) or (N,) when ``aligned``.  ``mode`` 'iou' | 'iof' (first
    argument is the foreground).  ``clockwise=False`` negates both angles,
    which leaves every value unchanged (mirror invariance), so it is accepted
    and has no effect.
    """
    return _overlaps(bboxes1, bboxes2, mode, aligned, 0.0)


_SEG_CAP = 16384     # segments one call handles (ids must also span fewer than 2**16 values)


def _nms_keep(boxes, scores, iou_threshold, seg_ids=None, offsets=None, id_offsets=False):
    """kept indices (int64, score-descending, on the CUDA device) of a greedy
    rotated NMS.

    ``seg_ids`` (ints per box) restricts suppression to boxes of equal id.  Centre
    offsets (added to cx, cy before the geometry, as the reference's class
    separation does): ``offsets`` (float per box), or ``id_offsets=True`` for
    mmcv.batched_nms's own ``id * (max(cx,cy) + max(w,h) + 1)``, computed on the
    device.  Two C-ABI calls: rotgeo_nms_prepare (key transform, radix sorts,
    segment sizes -> host: the one sync before the launch) and rotgeo_nms_prepared
    (pack + mask + resolve kernels), then one read of the kept count."""
    dev = _cuda_device(boxes, scores)
    n = boxes.size(0)
    b = _rows_f32(boxes, dev)
    s = scores.detach().to(device=dev, dtype=torch.float32).reshape(-1).contiguous()
    if not float(iou_threshold) > 0.0:
        return _top1(s)
    L = _lib.load()
    with torch.cuda.device(dev):
        seg_ptr = None
        if seg_ids is not None:
            seg = seg_ids.detach().to(device=dev, dtype=torch.int64).reshape(-1).contiguous()
            seg_ptr = seg.data_ptr()
        off_ptr = None
        if offsets is not None:
            offsets = offsets.detach().to(device=dev, dtype=torch.float32).contiguous()
            off_ptr = offsets.data_ptr()
        stream = _stream(dev)
        pbytes = L.rotgeo_nms_prepare_bytes(n)
        prep = torch.empty(pbytes, dtype=torch.uint8, device=dev)
        cap = min(n, _SEG_CAP)
        counts = np.empty(cap, dtype=np.int64)
        nseg = ctypes.c_int64(0)
        rc = L.rotgeo_nms_prepare(b.data_ptr(), n, b.stride(0), s.data_ptr(), seg_ptr, prep.data_ptr(),
                                  pbytes, counts.ctypes.data, cap, ctypes.byref(nseg), stream)
        if rc == _lib.ERR_RANGE:
            return _nms_keep_split(b, s, iou_threshold, seg, offsets, id_offsets)
        _lib.check(rc, 'rotgeo_nms_prepare')
        nseg = nseg.value
        seg_off = np.zeros(nseg + 1, dtype=np.int64)
        np.cumsum(counts[:nseg], out=seg_off[1:])
        nbytes = L.rotgeo_nms_workspace_bytes(n, seg_off.ctypes.data, nseg)
        ws = torch.empty(nbytes, dtype=torch.uint8, device=dev)
        keep = torch.empty(n, dtype=torch.int64, device=dev)
        nkeep = torch.empty(1, dtype=torch.int64, device=dev)
        _lib.check(L.rotgeo_nms_prepared(b.data_ptr(), n, b.stride(0), off_ptr, seg_ptr,
                                         1 if id_offsets else 0, prep.data_ptr(), counts.ctypes.data,
                                         nseg, float(iou_threshold), keep.data_ptr(), nkeep.data_ptr(),
                                         ws.data_ptr(), nbytes, stream), 'rotgeo_nms_prepared')
        k = int(nkeep.item())          # the one sync after the launch: sizes the output
    return keep[:k]


def _top1(s, lo=0, hi=None):
    """iou_threshold <= 0: the reference's CPU kernel tests ``IoU >= thr``, which holds for
    EVERY pair (IoU >= 0, boxes of other classes / levels included: their offset copies are
    disjoint, IoU 0) — the best-scored box suppresses all others.  The pair culls of the CUDA
    path never form disjoint pairs, so this case is answered here (ties: lower index)."""
    hi = s.numel() if hi is None else hi
    if hi <= lo:
        return torch.zeros(0, dtype=torch.int64, device=s.device)
    v = s[lo:hi]
    return torch.nonzero(v == v.max()).reshape(-1)[:1] + lo


def _nms_keep_images(boxes, scores, ids, sizes, iou_threshold):
    """Greedy rotated NMS of several images stored back to back (image g owns
    ``sizes[g]`` consecutive rows), one segment per (image, id), with
    mmcv.ops.batched_nms's id offsets computed per image on the device.

    Returns ``(keep, counts)``: ``keep`` (device int64) holds, image after image,
    the kept rows as indices INSIDE their image in descending score order;
    ``counts`` (python list) the number kept per image.  ``None`` when the
    (image, id) pairs do not fit 16-bit segment keys (callers use the generic
    path).  Three C-ABI calls (rotgeo_nms_batch_prepare: sorts + the segment
    sizes, one sync; rotgeo_nms_prepared: pack + mask + resolve;
    rotgeo_nms_batch_finish: per-image indices and counts) and one read of the
    counts."""
    dev = _cuda_device(boxes, scores)
    n = boxes.size(0)
    nimg = len(sizes)
    b = _rows_f32(boxes, dev)
    s = scores.detach().to(device=dev, dtype=torch.float32).reshape(-1).contiguous()
    seg = ids.detach().to(device=dev, dtype=torch.int64).reshape(-1).contiguous()
    if not float(iou_threshold) > 0.0:
        st = np.concatenate([[0], np.cumsum(sizes)])
        tops = [_top1(s, int(st[g]), int(st[g + 1])) - int(st[g]) for g in range(nimg)]
        return torch.cat(tops), [int(t.numel()) for t in tops]
    L = _lib.load()
    with torch.cuda.device(dev):
        stream = _stream(dev)
        sizes_h = np.ascontiguousarray(sizes, dtype=np.int64)
        pbytes = L.rotgeo_nms_batch_prepare_bytes(n, nimg)
        prep = torch.empty(pbytes, dtype=torch.uint8, device=dev)
        offsets = torch.empty(n, dtype=torch.float32, device=dev)
        cap = min(n, _SEG_CAP)
        counts = np.empty(cap, dtype=np.int64)
        nseg = ctypes.c_int64(0)
        rc = L.rotgeo_nms_batch_prepare(b.data_ptr(), n, b.stride(0), s.data_ptr(), seg.data_ptr(),
                                        sizes_h.ctypes.data, nimg, prep.data_ptr(), pbytes,
                                        offsets.data_ptr(), counts.ctypes.data, cap, ctypes.byref(nseg), stream)
        if rc == _lib.ERR_RANGE:
            return None
        _lib.check(rc, 'rotgeo_nms_batch_prepare')
        nseg = nseg.value
        seg_off = np.zeros(nseg + 1, dtype=np.int64)
        np.cumsum(counts[:nseg], out=seg_off[1:])
        nbytes = L.rotgeo_nms_workspace_bytes(n, seg_off.ctypes.data, nseg)
        ws = torch.empty(nbytes, dtype=torch.uint8, device=dev)
        keep = torch.empty(n, dtype=torch.int64, device=dev)
        tail = torch.empty(nimg + 1, dtype=torch.int64, device=dev)     # [kept per image ..., kept in all]
        _lib.check(L.rotgeo_nms_prepared(b.data_ptr(), n, b.stride(0), offsets.data_ptr(), seg.data_ptr(), 0,
                                         prep.data_ptr(), counts.ctypes.data, nseg, float(iou_threshold),
                                         keep.data_ptr(), tail.data_ptr() + 8 * nimg, ws.data_ptr(), nbytes,
                                         stream), 'rotgeo_nms_prepared')
        _lib.check(L.rotgeo_nms_batch_finish(keep.data_ptr(), tail.data_ptr() + 8 * nimg, n, prep.data_ptr(),
                                             nimg, tail.data_ptr(), stream), 'rotgeo_nms_batch_finish')
        per_img = tail.tolist()        # the one sync after the launches: sizes the outputs
    return keep[:per_img[-1]], per_img[:-1]


def _nms_keep_split(b, s, iou_threshold, seg, offsets, id_offsets):
    """segment ids too spread out / too many segments for one call: renumber them
    densely (torch.unique) and run chunks of _SEG_CAP segments; the centre offsets
    are materialised first so that every chunk sees the same values."""
    if id_offsets:
        max_coordinate = b[:, :2].max() + b[:, 2:4].max()
        offsets = seg.to(b.dtype) * (max_coordinate + torch.tensor(1).to(b))
    dense = torch.unique(seg, sorted=True, return_inverse=True)[1]
    nseg = int(dense.max().item()) + 1
    keeps = []
    for lo in range(0, nseg, _SEG_CAP):
        sel = torch.nonzero((dense >= lo) & (dense < lo + _SEG_CAP)).squeeze(1)
        k = _nms_keep(b[sel], s[sel], iou_threshold, seg_ids=dense[sel] - lo,
                      offsets=None if offsets is None else offsets[sel])
        keeps.append(sel[k])
    keep = torch.cat(keeps)
    # chunks -> global score order (ties: lower index first)
    order = torch.sort(keep).values
    return order[torch.sort(s[order], descending=True, stable=True).indices]


def nms_threshold_band(boxes, iou_threshold, seg_ids=None, offsets=None, tol=1e-5, chunk=4096):
    """Pairs whose IoU lies within ``tol`` of ``iou_threshold`` — the only pairs at which the
    kept set may legitimately differ from the reference's (BASELINE.json north_star: "keep
    indices identical except at pairs whose IoU lies within 1e-5 of the threshold, which are
    listed").  Returns ``(pairs (k, 2) int64 with i < j, ious (k,))`` on the CUDA device;
    pairs of different ``seg_ids`` are not listed.  ``offsets`` as in ``_nms_keep``.  Diagnostic
    helper: the (chunk x n) IoU blocks are materialised, sizes of tests and audits only."""
    dev = _cuda_device(boxes)
    b = _rows_f32(boxes, dev).clone()
    if offsets is not None:
        b[:, :2] += offsets.to(device=dev, dtype=torch.float32)[:, None]
    n = b.shape[0]
    seg = None if seg_ids is None else seg_ids.to(device=dev, dtype=torch.int64).reshape(-1)
    pairs, vals = [], []
    for lo in range(0, n, chunk):
        hi = min(n, lo + chunk)
        iou = _overlaps(b[lo:hi], b, 'iou', False, 0.0)
        near = (iou - float(iou_threshold)).abs() <= tol
        near &= torch.arange(lo, hi, device=dev)[:, None] < torch.arange(n, device=dev)[None, :]
        if seg is not None:
            near &= seg[lo:hi, None] == seg[None, :]
        ij = near.nonzero(as_tuple=False)
        if ij.numel():
            vals.append(iou[ij[:, 0], ij[:, 1]])
            ij[:, 0] += lo
            pairs.append(ij)
    if not pairs:
        return torch.zeros((0, 2), dtype=torch.int64, device=dev), torch.zeros(0, device=dev)
    return torch.cat(pairs), torch.cat(vals)


def nms_rotated(dets, scores, iou_threshold, labels=None, clockwise=True):
    """Drop-in for ``mmcv.ops.nms_rotated``.

    Returns ``(cat(dets[keep], scores[keep,None]) (K,6), keep (K,) int64)`` with
    ``keep`` indexing the input in descending-score order, and ``(dets, None)``
    for empty input.  Suppression follows the reference's CPU kernel
    (``IoU >= iou_threshold``).  ``labels``: mmcv appends them as a 6th column
    that neither of its kernels compares, so they do not separate classes there
    and do not here (the reference never passes them; class separation is the
    offset trick of bbox_nms_rotated.py:73-77).  ``clockwise`` has no effect on
    IoU (see box_iou_rotated).
    """
    if dets.shape[0] == 0:
        return dets, None
    assert dets.size(1) == 5, 'dets must be (N, 5) rotated boxes'
    keep = _nms_keep(dets, scores, iou_threshold)
    keep = keep.to(dets.device)
    out = torch.cat((dets[keep], scores[keep].reshape(-1, 1).to(dets.dtype)), dim=1)
    return out, keep


def _nms_horizontal(boxes, scores, iou_threshold, **kwargs):
    """axis-aligned ``mmcv.ops.nms`` stand-in (OUT OF north_star scope, library
    call: the fork's RPN configs select type='nms' on obb2xyxy boxes,
    configs/oriented_rcnn/oriented_rcnn_r50_fpn_1x_dota_le90.py:93,116)."""
    from torchvision.ops import nms as tv_nms
    dev = _cuda_device(boxes, scores)
    keep = tv_nms(boxes.to(dev).float(), scores.to(dev).float(), float(iou_threshold))
    keep = keep.to(boxes.device)
    return torch.cat((boxes[keep], scores[keep].reshape(-1, 1)), dim=1), keep


_NMS_TYPES = {'nms_rotated': nms_rotated, 'nms': _nms_horizontal}


def batched_nms(boxes, scores, idxs, nms_cfg, class_agnostic=False):
    """Drop-in for ``mmcv.ops.batched_nms`` (called at
    reference oriented_rpn_head.py:275, rotated_rpn_head.py:515).

    NMS among boxes of equal ``idxs`` only.  ``nms_cfg``: dict with ``type``
    ('nms_rotated' for 5-column boxes, 'nms' for 4-column ones),
    ``iou_threshold``/``iou_thr``, optional ``class_agnostic``, ``split_thr``,
    ``max_num``.  Returns ``(cat(boxes[keep], scores[keep,None]), keep)`` in
    descending-score order.  The rotated type runs as ONE segmented launch
    whatever the box count (mmcv loops over ids once N >= split_thr; the kept
    set and order are the same).
    """
    if nms_cfg is None:
        scores, inds = scores.sort(descending=True)
        return torch.cat([boxes[inds], scores[:, None]], -1), inds
    cfg = dict(nms_cfg)
    class_agnostic = cfg.pop('class_agnostic', class_agnostic)
    nms_type = cfg.pop('type', 'nms')
    cfg.pop('split_thr', None)
    max_num = cfg.pop('max_num', -1)
    if nms_type not in _NMS_TYPES:
        raise KeyError('unsupported nms type %r' % (nms_type,))
    thr = cfg.pop('iou_threshold', cfg.pop('iou_thr', None))
    if thr is None:
        raise TypeError("nms_cfg needs 'iou_threshold'")
    if boxes.shape[0] == 0:
        return torch.cat([boxes, scores[:, None]], -1), boxes.new_zeros((0,), dtype=torch.long)

    if nms_type == 'nms_rotated' and boxes.size(-1) == 5:
        # not class_agnostic: the centre offset  id * (max(cx,cy) + max(w,h) + 1)  of mmcv, same
        # fp32 arithmetic, derived on the device
        keep = _nms_keep(boxes, scores, thr, seg_ids=None if class_agnostic else idxs,
                         id_offsets=not class_agnostic).to(boxes.device)
    else:
        if class_agnostic:
            b = boxes
        else:
            max_coordinate = boxes.max()
            off = idxs.to(boxes) * (max_coordinate + torch.tensor(1).to(boxes))
            b = boxes + off[:, None]
        _, keep = _nms_horizontal(b, scores, thr)
    if max_num > 0:
        keep = keep[:max_num]
    out = torch.cat([boxes[keep], scores[keep][:, None]], -1)
    return out, keep
