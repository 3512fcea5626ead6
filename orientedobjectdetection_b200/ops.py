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
import contextlib
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


_NULL_CTX = contextlib.nullcontext()


def _on_device(dev):
    """``torch.cuda.device(dev)`` only when ``dev`` is not current already (the context manager costs
    ~10 us per entry, a tenth of a small NMS call)"""
    return _NULL_CTX if torch.cuda.current_device() == dev.index else torch.cuda.device(dev)


_RAW_STREAM = getattr(torch._C, '_cuda_getCurrentRawStream', None)


def _stream(dev):
    """the raw handle of torch's current stream on ``dev`` (torch.cuda.current_stream() builds a Stream object
    and costs ~7 us per call — 5 % of a small NMS call; the private accessor, which Triton's launcher uses too,
    costs a fraction of a microsecond)"""
    if _RAW_STREAM is not None:
        return _RAW_STREAM(dev.index)
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
    if not aligned and not bboxes1.is_cuda and rows * cols * 4 >= _HOST_CHUNK_BYTES and rows >= 4 * 64:
        out = _overlaps_to_host_chunked(L, dev, b1, b2, rows, cols, mode, min_wh)
        return out if out.dtype == bboxes1.dtype else out.to(dtype=bboxes1.dtype)
    with _on_device(dev):
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


_HOST_CHUNK_BYTES = 32 << 20
_COPY_STREAMS = {}


def _overlaps_to_host_chunked(L, dev, b1, b2, rows, cols, mode, min_wh):
    """A big matrix for a CPU caller is bound by the device->host copy (268 MB at ~57 GB/s = 4.7 ms
    against 0.05 ms of kernel at 8192 x 8192): four row chunks, each one rotgeo_iou_pairwise call on
    the compute stream, its copy into the pinned result issued on a second stream as soon as the
    chunk is done — the copies run back to back from the first quarter of the kernel time on,
    instead of after the whole matrix."""
    with _on_device(dev):
        compute = torch.cuda.current_stream(dev)
        copy = _COPY_STREAMS.get(dev.index)
        if copy is None:
            copy = _COPY_STREAMS[dev.index] = torch.cuda.Stream(dev)
        out = torch.empty((rows, cols), dtype=torch.float32, device=dev)
        out.record_stream(copy)
        host = torch.empty((rows, cols), dtype=torch.float32, pin_memory=True)
        step = (rows + 3) // 4
        step = (step + 63) // 64 * 64
        nbytes = L.rotgeo_iou_workspace_bytes(step, cols)
        ws = _workspace(dev, compute.cuda_stream, 'iou_host', nbytes)
        for lo in range(0, rows, step):
            hi = min(rows, lo + step)
            _lib.check(L.rotgeo_iou_pairwise(b1.data_ptr() + lo * b1.stride(0) * 4, hi - lo, b1.stride(0), b2.data_ptr(),
                                             cols, b2.stride(0), _MODE[mode], float(min_wh),
                                             out.data_ptr() + lo * cols * 4, cols, ws.data_ptr(), ws.numel(),
                                             compute.cuda_stream), 'rotgeo_iou_pairwise')
            ev = torch.cuda.Event()
            ev.record(compute)
            copy.wait_event(ev)
            with torch.cuda.stream(copy):
                host[lo:hi].copy_(out[lo:hi], non_blocking=True)
        copy.synchronize()
    return host


def box_iou_rotated(bboxes1, bboxes2, mode='iou', aligned=False, clockwise=True):
    """Drop-in for ``mmcv.ops.box_iou_rotated``.

    ``bboxes1`` (N,5+) and ``bboxes2`` (M,5+) rows ``(cx, cy, w, h, theta)``;
    returns (N, M) or (N,) when ``aligned``.  ``mode`` 'iou' | 'iof' (first
    argument is the foreground).  ``clockwise=False`` negates both angles,
    which leaves every value unchanged (mirror invariance), so it is accepted
    and has no effect.
    """
    return _overlaps(bboxes1, bboxes2, mode, aligned, 0.0)


_SEG_CAP = 16384          # segments one call handles (the ids of a group must also span fewer than 2**16 values)
_PLAN_LIMIT = 4 << 30     # worst-case workspace above this: read the device plan back once instead (the
                          # workspace is cached per stream, so its size costs memory, not time)
_ERR_RANGE_BIT, _ERR_SEGMENTS_BIT = 1, 2

_WS = {}


def _workspace(dev, stream, name, nbytes):
    """Reusable device scratch per (device, stream, purpose): grows, never shrinks.  The C-ABI
    takes caller-owned workspace; allocating it per call cost more than small NMS calls take."""
    key = (dev.index, stream, name)
    t = _WS.get(key)
    if t is None or t.numel() < nbytes:
        t = torch.empty(max(int(nbytes * 1.25), 1 << 16), dtype=torch.uint8, device=dev)
        _WS[key] = t
    return t


def release_workspaces():
    """drop the cached scratch buffers (they are plain torch tensors)"""
    _WS.clear()


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


def _nms_grouped(boxes, scores, iou_threshold, seg_ids=None, offsets=None, id_offsets=False,
                 group_sizes=None, seg_cap=None, no_read=False, with_dets=False):
    """Greedy rotated NMS of one or several GROUPS of boxes stored back to back (group g owns
    ``group_sizes[g]`` consecutive rows; None = one group).  Suppression only among boxes of
    equal ``seg_ids`` inside a group.  Centre offsets (added to cx, cy before the geometry, as
    the reference's class separation does): ``offsets`` (float per box), or ``id_offsets=True``
    for mmcv.batched_nms's own ``id * (max(cx,cy) + max(w,h) + 1)`` with the maxima of the
    box's own group, computed on the device.

    Returns ``(keep, counts)``: ``keep`` (device int64) holds, group after group, the kept rows
    in descending score order — global row indices when ``group_sizes`` is None, else indices
    INSIDE the group; ``counts`` the kept boxes per group (python list, None for one group).
    ``None`` when the ids do not fit the device front-end (callers renumber and split).

    Launch sequence (C-ABI): rotgeo_nms_sort (ONE kernel: orders, offsets, segment table and
    launch totals in a device header) and rotgeo_nms_run (pack + mask + resolve over worst-case
    bounds): no device->host read until the kept count is fetched at the end.  Jobs whose
    worst-case workspace exceeds _PLAN_LIMIT read the 64-byte plan back once instead.

    ``with_dets=True``: returns ``(keep, counts, dets)`` with ``dets`` (k, 6) float32 =
    ``cat(boxes[keep], scores[keep, None])`` written by the resolve kernel itself (None when the call took
    the plan-read path; the caller gathers then).

    ``no_read=True`` (the multi-GPU callers, which push the result to their peers first and read
    once at the very end): nothing is read back; returns ``(keep (n,), tail)`` with ``tail`` the
    device int64 vector ``[kept in all, front-end error bits, kept per group ...]``; a non-zero
    error word means the result is void (ids the front-end does not take)."""
    dev = _cuda_device(boxes, scores)
    n = boxes.size(0)
    b = _rows_f32(boxes, dev)
    s = scores.detach().to(device=dev, dtype=torch.float32).reshape(-1).contiguous()
    grouped = group_sizes is not None
    sizes_h = np.ascontiguousarray(group_sizes, dtype=np.int64) if grouped else None
    ngroups = int(sizes_h.shape[0]) if grouped else 1
    if no_read:
        assert grouped and n > 0 and float(iou_threshold) > 0.0
    if with_dets and (n == 0 or not float(iou_threshold) > 0.0):
        got = _nms_grouped(boxes, scores, iou_threshold, seg_ids, offsets, id_offsets, group_sizes, seg_cap)
        return got[0], got[1], None
    if n == 0:
        return torch.zeros(0, dtype=torch.int64, device=dev), ([0] * ngroups if grouped else None)
    if not float(iou_threshold) > 0.0:
        if not grouped:
            return _top1(s), None
        st = np.concatenate([[0], np.cumsum(sizes_h)])
        tops = [_top1(s, int(st[g]), int(st[g + 1])) - int(st[g]) for g in range(ngroups)]
        return torch.cat(tops), [int(t.numel()) for t in tops]
    L = _lib.load()
    with _on_device(dev):
        seg_ptr = None
        if seg_ids is not None:
            seg = seg_ids.detach().to(device=dev, dtype=torch.int64).reshape(-1).contiguous()
            seg_ptr = seg.data_ptr()
        off_ptr = None
        if offsets is not None:
            offsets = offsets.detach().to(device=dev, dtype=torch.float32).contiguous()
            off_ptr = offsets.data_ptr()
        stream = _stream(dev)
        sizes_ptr = sizes_h.ctypes.data if grouped else None
        max_group = int(sizes_h.max()) if grouped else n
        if seg_ptr is None:
            cap = 1
        elif seg_cap is not None:
            cap = int(seg_cap)
        else:
            cap = min(max(max_group, 1), 256 if ngroups == 1 else 16, _SEG_CAP)
        use_sort_off = 1 if (id_offsets and seg_ptr) else 0
        sizes = np.zeros(6, dtype=np.int64)
        need = (ctypes.c_size_t * 2)()
        for attempt in range(2):
            ints = torch.empty(n + 2 + ngroups, dtype=torch.int64, device=dev)   # one allocation: keep | tail
            keep, tail = ints[:n], ints[n:]                                      # tail = kept, error bits, kept per group
            dets = torch.empty((n, 6), dtype=torch.float32, device=dev) if with_dets else None
            # fast path: ONE C-ABI call (sort + bounds + run) over the cached buffers; it reports the sizes it
            # needs, so a buffer that is too small costs one retry, once
            launched = False
            skey_s, skey_w = (dev.index, stream, 'nms_sort'), (dev.index, stream, 'nms_run')
            for _ in range(2):
                sortws, ws = _WS.get(skey_s), _WS.get(skey_w)
                rc = L.rotgeo_nms_grouped(b.data_ptr(), n, b.stride(0), s.data_ptr(), seg_ptr, use_sort_off, off_ptr,
                                          sizes_ptr, ngroups, cap, float(iou_threshold), keep.data_ptr(),
                                          tail.data_ptr(), 1 if grouped else 0,
                                          dets.data_ptr() if dets is not None else None,
                                          sortws.data_ptr() if sortws is not None else None,
                                          sortws.numel() if sortws is not None else 0,
                                          ws.data_ptr() if ws is not None else None,
                                          ws.numel() if ws is not None else 0, need, stream)
                if rc == _lib.ERR_WORKSPACE and need[1] <= _PLAN_LIMIT:
                    _workspace(dev, stream, 'nms_sort', need[0])
                    _workspace(dev, stream, 'nms_run', need[1])
                    continue
                if rc != _lib.ERR_WORKSPACE:
                    _lib.check(rc, 'rotgeo_nms_grouped')
                    launched = True
                break
            err = 0
            if not launched:
                dets = None
                # worst-case workspace not worth allocating: sort, read the 64-byte plan back once, run exactly
                sbytes = L.rotgeo_nms_sort_bytes(n, sizes_ptr, ngroups, cap)
                sortws = _workspace(dev, stream, 'nms_sort', sbytes)
                _lib.check(L.rotgeo_nms_sort(b.data_ptr(), n, b.stride(0), s.data_ptr(), seg_ptr, use_sort_off,
                                             sizes_ptr, ngroups, cap, sortws.data_ptr(), sortws.numel(), stream),
                           'rotgeo_nms_sort')
                rc = L.rotgeo_nms_read_plan(sortws.data_ptr(), n, sizes_ptr, ngroups, cap, sizes.ctypes.data, stream)
                if rc == _lib.ERR_RANGE:
                    err = _plan_error_bits()
                else:
                    _lib.check(rc, 'rotgeo_nms_read_plan')
                    nbytes = L.rotgeo_nms_run_bytes(n, sizes.ctypes.data)
                    ws = _workspace(dev, stream, 'nms_run', nbytes)
                    _lib.check(L.rotgeo_nms_run(b.data_ptr(), n, b.stride(0), off_ptr, use_sort_off, sizes_ptr, ngroups,
                                                cap, sortws.data_ptr(), sizes.ctypes.data, float(iou_threshold),
                                                keep.data_ptr(), tail.data_ptr(),
                                                tail.data_ptr() + 16 if grouped else None,
                                                ws.data_ptr(), ws.numel(), stream), 'rotgeo_nms_run')
            if not err:
                if no_read:
                    return keep, tail
                host = tail.tolist()          # the one device->host read of the call: sizes the outputs
                err = int(host[1])
                if not err:
                    if with_dets:
                        return keep[:host[0]], (host[2:] if grouped else None), (dets[:host[0]] if dets is not None else None)
                    return keep[:host[0]], (host[2:] if grouped else None)
            if (err & _ERR_RANGE_BIT) or attempt == 1 or cap >= min(max_group, _SEG_CAP):
                return None
            cap = min(max_group, _SEG_CAP)     # more segments in a group than assumed: once more with room
    return None


def _plan_error_bits():
    msg = _lib.load().rotgeo_last_error().decode('utf-8', 'replace')
    return _ERR_RANGE_BIT if 'span' in msg else _ERR_SEGMENTS_BIT


def _nms_keep(boxes, scores, iou_threshold, seg_ids=None, offsets=None, id_offsets=False, with_dets=False):
    """kept indices (int64, score-descending, on the CUDA device) of a greedy rotated NMS of
    ONE group (see _nms_grouped); ids that the device front-end does not take (spread over
    65536 values or more, more than 16384 segments) are renumbered and split.
    ``with_dets``: returns ``(keep, dets (k, 6) float32 | None)``."""
    got = _nms_grouped(boxes, scores, iou_threshold, seg_ids, offsets, id_offsets, with_dets=with_dets)
    if got is not None:
        return (got[0], got[2]) if with_dets else got[0]
    if with_dets:
        return _nms_keep(boxes, scores, iou_threshold, seg_ids, offsets, id_offsets), None
    dev = _cuda_device(boxes, scores)
    b = _rows_f32(boxes, dev)
    s = scores.detach().to(device=dev, dtype=torch.float32).reshape(-1).contiguous()
    seg = seg_ids.detach().to(device=dev, dtype=torch.int64).reshape(-1)
    if offsets is not None:
        offsets = offsets.detach().to(device=dev, dtype=torch.float32)
    return _nms_keep_split(b, s, iou_threshold, seg, offsets, id_offsets)


def _nms_keep_images(boxes, scores, ids, sizes, iou_threshold):
    """Greedy rotated NMS of several images stored back to back (image g owns
    ``sizes[g]`` consecutive rows), one segment per (image, id), with
    mmcv.ops.batched_nms's id offsets computed per image on the device.

    Returns ``(keep, counts)``: ``keep`` (device int64) holds, image after image,
    the kept rows as indices INSIDE their image in descending score order;
    ``counts`` (python list) the number kept per image.  ``None`` when the ids of
    an image do not fit the device front-end (callers use the generic path)."""
    return _nms_grouped(boxes, scores, iou_threshold, seg_ids=ids, id_offsets=True, group_sizes=sizes)


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
        got = _nms_grouped(b[sel], s[sel], iou_threshold, seg_ids=dense[sel] - lo,
                           offsets=None if offsets is None else offsets[sel], seg_cap=_SEG_CAP)
        if got is None:
            raise RuntimeError('rotgeo: the NMS front-end rejected a renumbered job')
        keeps.append(sel[got[0]])
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
    fused = dets.is_cuda and dets.dtype == torch.float32 and scores.dtype == torch.float32
    keep, out = _nms_keep(dets, scores, iou_threshold, with_dets=fused) if fused else (_nms_keep(dets, scores, iou_threshold), None)
    keep = keep.to(dets.device)
    if out is None:      # other dtypes / host tensors: the caller's own values, gathered
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
        fused = boxes.is_cuda and boxes.dtype == torch.float32 and scores.dtype == torch.float32 and max_num <= 0
        if fused:
            keep, out = _nms_keep(boxes, scores, thr, seg_ids=None if class_agnostic else idxs,
                                  id_offsets=not class_agnostic, with_dets=True)
            if out is not None:
                return out, keep
        else:
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
