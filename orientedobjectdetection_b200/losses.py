"""Differentiable rotated IoU and the IoU loss built on it.

``diff_iou_rotated_2d`` has the signature of the mmcv op of that name as the reference's
loss calls it (reference mmrotate/models/losses/rotated_iou_loss.py:10,47): boxes
``(B, N, 5)`` as ``(x, y, w, h, alpha)`` -> IoU ``(B, N)``, differentiable with respect to
both inputs.  Value and gradient come from ONE kernel (rotgeo_diff_iou_aligned; analytic
gradient, see csrc/rotgeo_diff.cu) instead of mmcv's chain of torch ops + vertex sort +
autograd.  ``rotated_iou_loss`` / ``RotatedIoULoss`` mirror
rotated_iou_loss.py:16-58 / 61-140 with mmdet's ``weighted_loss`` reduction restated
(mmdet/models/losses/utils.py, external dependency of the reference).  No CPU fallback.
"""
import warnings

import torch
import torch.nn as nn

from . import _lib
from .ops import _cuda_device, _on_device, _stream


class _DiffIoU(torch.autograd.Function):
    @staticmethod
    def forward(ctx, box1, box2):
        dev = _cuda_device(box1, box2)
        shape = box1.shape[:-1]
        a = box1.detach().reshape(-1, 5).to(device=dev, dtype=torch.float32).contiguous()
        b = box2.detach().reshape(-1, 5).to(device=dev, dtype=torch.float32).contiguous()
        n = a.shape[0]
        L = _lib.load()
        with _on_device(dev):
            iou = torch.empty(n, dtype=torch.float32, device=dev)
            need1, need2 = ctx.needs_input_grad[0], ctx.needs_input_grad[1]
            g1 = torch.empty((n, 5), dtype=torch.float32, device=dev) if need1 else None
            g2 = torch.empty((n, 5), dtype=torch.float32, device=dev) if need2 else None
            _lib.check(L.rotgeo_diff_iou_aligned(a.data_ptr(), b.data_ptr(), n, 5, 5, iou.data_ptr(),
                                                 g1.data_ptr() if need1 else None,
                                                 g2.data_ptr() if need2 else None, _stream(dev)),
                       'rotgeo_diff_iou_aligned')
        ctx.save_for_backward(*(t for t in (g1, g2) if t is not None))
        ctx.flags = (need1, need2)
        ctx.meta = (box1.shape, box1.dtype, box1.device, box2.shape, box2.dtype, box2.device)
        return iou.reshape(shape).to(device=box1.device, dtype=box1.dtype)

    @staticmethod
    def backward(ctx, grad_out):
        saved = list(ctx.saved_tensors)
        need1, need2 = ctx.flags
        s1, t1, d1, s2, t2, d2 = ctx.meta
        out = [None, None]
        if need1:
            g = saved.pop(0)
            out[0] = (g * grad_out.reshape(-1, 1).to(g)).reshape(s1).to(device=d1, dtype=t1)
        if need2:
            g = saved.pop(0)
            out[1] = (g * grad_out.reshape(-1, 1).to(g)).reshape(s2).to(device=d2, dtype=t2)
        return tuple(out)


def diff_iou_rotated_2d(box1, box2):
    """mmcv.ops.diff_iou_rotated_2d: ``box1``, ``box2`` (B, N, 5) -> IoU (B, N)."""
    assert box1.shape == box2.shape and box1.size(-1) == 5
    return _DiffIoU.apply(box1, box2)


def _weight_reduce_loss(loss, weight=None, reduction='mean', avg_factor=None):
    """mmdet.models.losses.utils.weight_reduce_loss"""
    if weight is not None:
        loss = loss * weight
    if avg_factor is None:
        if reduction == 'mean':
            return loss.mean()
        if reduction == 'sum':
            return loss.sum()
        return loss
    if reduction == 'mean':
        eps = torch.finfo(torch.float32).eps
        return loss.sum() / (avg_factor + eps)
    if reduction != 'none':
        raise ValueError('avg_factor can not be used with reduction="sum"')
    return loss


def rotated_iou_loss(pred, target, weight=None, linear=False, mode='log', eps=1e-6, reduction='mean',
                     avg_factor=None):
    """reference rotated_iou_loss.py:16-58 including the ``@weighted_loss`` wrapper"""
    assert mode in ['linear', 'square', 'log']
    if linear:
        mode = 'linear'
        warnings.warn('DeprecationWarning: Setting "linear=True" in poly_iou_loss is deprecated, '
                      'please use "mode=`linear`" instead.')
    ious = diff_iou_rotated_2d(pred.unsqueeze(0), target.unsqueeze(0))
    ious = ious.squeeze(0).clamp(min=eps)
    if mode == 'linear':
        loss = 1 - ious
    elif mode == 'square':
        loss = 1 - ious ** 2
    else:
        loss = -ious.log()
    return _weight_reduce_loss(loss, weight, reduction, avg_factor)


class RotatedIoULoss(nn.Module):
    """reference rotated_iou_loss.py:61-140 (same arguments, same forward)"""

    def __init__(self, linear=False, eps=1e-6, reduction='mean', loss_weight=1.0, mode='log'):
        super(RotatedIoULoss, self).__init__()
        assert mode in ['linear', 'square', 'log']
        if linear:
            mode = 'linear'
            warnings.warn('DeprecationWarning: Setting "linear=True" in IOULoss is deprecated, '
                          'please use "mode=`linear`" instead.')
        self.mode = mode
        self.linear = linear
        self.eps = eps
        self.reduction = reduction
        self.loss_weight = loss_weight

    def forward(self, pred, target, weight=None, avg_factor=None, reduction_override=None, **kwargs):
        assert reduction_override in (None, 'none', 'mean', 'sum')
        reduction = reduction_override if reduction_override else self.reduction
        if (weight is not None) and (not torch.any(weight > 0)) and (reduction != 'none'):
            if pred.dim() == weight.dim() + 1:
                weight = weight.unsqueeze(1)
            return (pred * weight).sum()  # 0
        if weight is not None and weight.dim() > 1:
            assert weight.shape == pred.shape
            weight = weight.mean(-1)
        return self.loss_weight * rotated_iou_loss(pred, target, weight, mode=self.mode, eps=self.eps,
                                                   reduction=reduction, avg_factor=avg_factor, **kwargs)
