"""Fused assigner reductions: what the reference's assigners do
with the matrix returned by ``RBboxOverlaps2D`` (reference
mmrotate/core/bbox/iou_calculators/rotate_iou2d_calculator.py:44), computed without the
matrix.

* ``rbbox_overlaps_max`` — the two ``overlaps.max`` reductions every assigner starts with
  (mmdet ``MaxIoUAssigner.assign_wrt_overlaps``; reference
  mmrotate/core/bbox/assigners/atss_obb_assigner.py:118-145, obb_label_assigner.py:128-143).
* ``MaxIoUAssigner`` — same constructor and ``assign`` signature as mmdet's class of that
  name, the assigner of the fork's RetinaNet / RPN / R-CNN configs (e.g. reference
  configs/rotated_retinanet/rotated_retinanet_obb_r50_fpn_1x_dota_le90.py:41-48 with
  ``iou_calculator=dict(type='RBboxOverlaps2D')``).  mmdet is an external dependency of the
  reference and is not installed here, so the result is a plain ``AssignResult`` namedtuple
  with mmdet's four fields.

Everything runs in librotgeo.so (rotgeo_iou_rowcol_max / rotgeo_max_iou_assign); there is
no CPU fallback.
"""
from collections import namedtuple

import torch

from . import _lib
from .ops import _MODE, _cuda_device, _on_device, _rows_f32, _stream

AssignResult = namedtuple('AssignResult', ['num_gts', 'gt_inds', 'max_overlaps', 'labels'])


def rbbox_overlaps_max(bboxes1, bboxes2, mode='iou', min_wh=1e-3):
    """``(row_max, row_argmax, col_max, col_argmax)`` of
    ``rbbox_overlaps(bboxes1, bboxes2, mode)`` — ``overlaps.max(dim=1)`` and
    ``overlaps.max(dim=0)`` — without materialising the (m, n) matrix.  The w, h clamp of
    rbbox_overlaps (rotate_iou2d_calculator.py:84-87) is applied (``min_wh``).  Lines without a
    positive value report 0 at index 0."""
    assert mode in _MODE
    assert bboxes1.size(-1) == 5 or bboxes1.size(0) == 0
    assert bboxes2.size(-1) == 5 or bboxes2.size(0) == 0
    dev = _cuda_device(bboxes1, bboxes2)
    n1, n2 = bboxes1.size(0), bboxes2.size(0)
    b1, b2 = _rows_f32(bboxes1, dev), _rows_f32(bboxes2, dev)
    L = _lib.load()
    with _on_device(dev):
        row_max = torch.empty(n1, dtype=torch.float32, device=dev)
        row_arg = torch.empty(n1, dtype=torch.int64, device=dev)
        col_max = torch.empty(n2, dtype=torch.float32, device=dev)
        col_arg = torch.empty(n2, dtype=torch.int64, device=dev)
        nbytes = L.rotgeo_iou_max_workspace_bytes(n1, n2)
        ws = torch.empty(max(nbytes, 16), dtype=torch.uint8, device=dev)
        _lib.check(L.rotgeo_iou_rowcol_max(b1.data_ptr(), n1, b1.stride(0) if n1 else 5, b2.data_ptr(), n2,
                                           b2.stride(0) if n2 else 5, _MODE[mode], float(min_wh),
                                           row_max.data_ptr(), row_arg.data_ptr(), col_max.data_ptr(),
                                           col_arg.data_ptr(), ws.data_ptr(), nbytes, _stream(dev)),
                   'rotgeo_iou_rowcol_max')
    out_dev = bboxes1.device
    return tuple(t.to(out_dev) for t in (row_max, row_arg, col_max, col_arg))


class MaxIoUAssigner(object):
    """mmdet ``MaxIoUAssigner`` with ``iou_calculator=dict(type='RBboxOverlaps2D')``, fused.

    Same arguments and semantics as mmdet 2.x ``max_iou_assigner.py`` for the paths the
    rotated configs of the reference use: ``assign(bboxes, gt_bboxes, gt_bboxes_ignore=None,
    gt_labels=None)`` returns gt_inds (-1 ignore, 0 background, i+1 = gt i), max_overlaps and
    labels.  ``gt_bboxes_ignore`` with ``ignore_iof_thr > 0`` and ``gpu_assign_thr`` are not
    supported (no rotated config of the reference sets them) and raise."""

    def __init__(self, pos_iou_thr, neg_iou_thr, min_pos_iou=.0, gt_max_assign_all=True,
                 ignore_iof_thr=-1, ignore_wrt_candidates=True, match_low_quality=True,
                 gpu_assign_thr=-1, iou_calculator=None):
        self.pos_iou_thr = pos_iou_thr
        self.neg_iou_thr = neg_iou_thr
        self.min_pos_iou = min_pos_iou
        self.gt_max_assign_all = gt_max_assign_all
        self.ignore_iof_thr = ignore_iof_thr
        self.ignore_wrt_candidates = ignore_wrt_candidates
        self.match_low_quality = match_low_quality
        self.gpu_assign_thr = gpu_assign_thr
        if iou_calculator is not None and dict(iou_calculator).get('type', 'RBboxOverlaps2D') != 'RBboxOverlaps2D':
            raise NotImplementedError('only the rotated IoU calculator is fused')

    def assign(self, bboxes, gt_bboxes, gt_bboxes_ignore=None, gt_labels=None):
        if (self.ignore_iof_thr > 0 and gt_bboxes_ignore is not None and gt_bboxes_ignore.numel() > 0
                and bboxes.numel() > 0):
            raise NotImplementedError('gt_bboxes_ignore is outside the fused path')
        if bboxes.size(-1) == 6:
            bboxes = bboxes[..., :5]
        if gt_bboxes.size(-1) == 6:
            gt_bboxes = gt_bboxes[..., :5]
        num_gts, num_bboxes = gt_bboxes.size(0), bboxes.size(0)
        if isinstance(self.neg_iou_thr, (tuple, list)):
            assert len(self.neg_iou_thr) == 2
            neg_lo, neg_hi = float(self.neg_iou_thr[0]), float(self.neg_iou_thr[1])
        else:
            neg_lo, neg_hi = 0.0, float(self.neg_iou_thr)
        out_dev = bboxes.device
        if num_bboxes == 0:
            gt_inds = bboxes.new_zeros((0,), dtype=torch.long)
            max_overlaps = bboxes.new_zeros((0,))
        else:
            dev = _cuda_device(bboxes, gt_bboxes)
            g, b = _rows_f32(gt_bboxes, dev), _rows_f32(bboxes, dev)
            L = _lib.load()
            with _on_device(dev):
                gt_inds = torch.empty(num_bboxes, dtype=torch.int64, device=dev)
                max_overlaps = torch.empty(num_bboxes, dtype=torch.float32, device=dev)
                nbytes = L.rotgeo_iou_max_workspace_bytes(num_gts, num_bboxes)
                ws = torch.empty(max(nbytes, 16), dtype=torch.uint8, device=dev)
                _lib.check(L.rotgeo_max_iou_assign(
                    g.data_ptr(), num_gts, g.stride(0) if num_gts else 5, b.data_ptr(), num_bboxes,
                    b.stride(0), 0, 1e-3, float(self.pos_iou_thr), neg_lo, neg_hi, float(self.min_pos_iou),
                    1 if self.match_low_quality else 0, 1 if self.gt_max_assign_all else 0,
                    gt_inds.data_ptr(), max_overlaps.data_ptr(), ws.data_ptr(), nbytes, _stream(dev)),
                    'rotgeo_max_iou_assign')
            gt_inds, max_overlaps = gt_inds.to(out_dev), max_overlaps.to(out_dev)
        labels = None
        if gt_labels is not None:
            labels = gt_inds.new_full((num_bboxes,), -1)
            pos = torch.nonzero(gt_inds > 0, as_tuple=False).squeeze(1)
            if pos.numel() > 0:
                labels[pos] = gt_labels.to(out_dev)[gt_inds[pos] - 1]
        return AssignResult(num_gts, gt_inds, max_overlaps, labels)
