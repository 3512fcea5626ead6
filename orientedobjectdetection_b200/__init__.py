"""rotgeo: B200-native (sm_100a) rotated-box IoU and rotated NMS behind the
``mmcv.ops`` / ``mmrotate.core`` call signatures of the MMRotate fork
goodmayonnaise/OrientedObjectDetection.

Importing the package touches neither CUDA nor the shared library; the first
operator call loads ``librotgeo.so`` and fails loudly if it is missing.
"""
from .core import (RBboxOverlaps2D, aug_multiclass_nms_rotated, merge_results,
                   multiclass_nms_rotated, rbbox_overlaps, translate_bboxes)
from .ops import batched_nms, box_iou_rotated, nms_rotated
from .assign import AssignResult, MaxIoUAssigner, rbbox_overlaps_max

__all__ = [
    'box_iou_rotated', 'nms_rotated', 'batched_nms', 'rbbox_overlaps', 'RBboxOverlaps2D',
    'multiclass_nms_rotated', 'aug_multiclass_nms_rotated', 'merge_results', 'translate_bboxes',
    'rbbox_overlaps_max', 'MaxIoUAssigner', 'AssignResult',
]
__version__ = '0.1.0'
