"""``patch()`` rebinds the reference's names to this package when mmcv /
mmrotate are importable (they are not in this image).  The
nine modules that bind the names at import time are listed in SURVEY.md §8(b).

Two reference call sites run in WORKER PROCESSES and are deliberately left on
the reference's own CPU op: ``RRandomCrop._crop_data`` (transforms.py:369-371,
DataLoader workers) and ``_merge_func`` (dota.py:358-382, reached through
``mmcv.track_parallel_progress(..., nproc=4)`` at dota.py:265-273).  A forked
child of a parent that has initialised CUDA cannot use the GPU, and this
package has no CPU fallback; the reference guards the second site with
``try: .cuda() / except: CPU`` (dota.py:374-377).  Instead,
``DOTADataset.merge_det`` itself is rebound to the single-process version that
merges every (image, class) group in one segmented launch, so ``_merge_func``
is no longer reached from it.
"""
import importlib

from . import assign, core, datasets, evaluation, losses, ops

_L1 = {'box_iou_rotated': ops.box_iou_rotated, 'nms_rotated': ops.nms_rotated,
       'batched_nms': ops.batched_nms, 'diff_iou_rotated_2d': losses.diff_iou_rotated_2d}
_L1_MODULES = (
    'mmcv.ops',
    'mmrotate.core.bbox.iou_calculators.rotate_iou2d_calculator',
    'mmrotate.core.evaluation.eval_map',
    # not 'mmrotate.datasets.pipelines.transforms': RRandomCrop runs in DataLoader workers
    'mmrotate.core.post_processing.bbox_nms_rotated',
    'mmrotate.core.patch.merge_results',
    'mmrotate.datasets.dota',
    'mmrotate.models.dense_heads.rotated_rpn_head',
    'mmrotate.models.dense_heads.oriented_rpn_head',
    'mmrotate.models.losses.rotated_iou_loss',
)
_L2 = {
    'mmrotate.core.bbox.iou_calculators.rotate_iou2d_calculator':
        {'rbbox_overlaps': core.rbbox_overlaps},
    'mmrotate.core.bbox.iou_calculators': {'rbbox_overlaps': core.rbbox_overlaps},
    'mmrotate.core.bbox': {'rbbox_overlaps': core.rbbox_overlaps},
    'mmrotate.core.post_processing.bbox_nms_rotated':
        {'multiclass_nms_rotated': core.multiclass_nms_rotated,
         'aug_multiclass_nms_rotated': core.aug_multiclass_nms_rotated},
    'mmrotate.core.post_processing':
        {'multiclass_nms_rotated': core.multiclass_nms_rotated,
         'aug_multiclass_nms_rotated': core.aug_multiclass_nms_rotated},
    'mmrotate.core.patch.merge_results': {'merge_results': core.merge_results},
    'mmrotate.core.patch': {'merge_results': core.merge_results},
    'mmrotate.core.evaluation.eval_map': {'eval_rbbox_map': evaluation.eval_rbbox_map},
    'mmrotate.core.evaluation': {'eval_rbbox_map': evaluation.eval_rbbox_map},
    # not dota._merge_func (pool workers, see above); DOTADataset.merge_det is rebound in patch()
    'mmrotate.datasets.dota': {'eval_rbbox_map': evaluation.eval_rbbox_map},
    'mmrotate.datasets.hrsc': {'eval_rbbox_map': evaluation.eval_rbbox_map},
    'mmrotate.core': {'rbbox_overlaps': core.rbbox_overlaps,
                      'multiclass_nms_rotated': core.multiclass_nms_rotated,
                      'aug_multiclass_nms_rotated': core.aug_multiclass_nms_rotated,
                      'merge_results': core.merge_results,
                      'eval_rbbox_map': evaluation.eval_rbbox_map},
}


def patch(strict=False):
    """Rebind every name; returns the list of (module, name) pairs patched.
    Modules that are not importable are skipped unless ``strict``."""
    done = []
    for modname in _L1_MODULES:
        try:
            mod = importlib.import_module(modname)
        except Exception:
            if strict:
                raise
            continue
        for name, fn in _L1.items():
            if hasattr(mod, name):
                setattr(mod, name, fn)
                done.append((modname, name))
    for modname, names in _L2.items():
        try:
            mod = importlib.import_module(modname)
        except Exception:
            if strict:
                raise
            continue
        for name, fn in names.items():
            if hasattr(mod, name):
                setattr(mod, name, fn)
                done.append((modname, name))
    try:   # the submission path: one segmented launch in THIS process instead of a pool of forked workers
        dota = importlib.import_module('mmrotate.datasets.dota')

        def merge_det(self, results, nproc=4):
            return datasets.merge_det(results, self.img_ids, self.CLASSES, nproc)   # (id_list, dets_list)

        dota.DOTADataset.merge_det = merge_det
        done.append(('mmrotate.datasets.dota', 'DOTADataset.merge_det'))
    except Exception:
        if strict:
            raise
    try:   # configs say iou_calculator=dict(type='RBboxOverlaps2D')
        from mmdet.core.bbox.iou_calculators import IOU_CALCULATORS
        IOU_CALCULATORS.register_module(name='RBboxOverlaps2D', force=True,
                                        module=core.RBboxOverlaps2D)
        done.append(('mmdet IOU_CALCULATORS', 'RBboxOverlaps2D'))
    except Exception:
        if strict:
            raise
    try:   # opt-in by name: assigner=dict(type='FusedMaxIoUAssigner', ...) in a config
        from mmdet.core.bbox.builder import BBOX_ASSIGNERS
        BBOX_ASSIGNERS.register_module(name='FusedMaxIoUAssigner', force=True, module=assign.MaxIoUAssigner)
        done.append(('mmdet BBOX_ASSIGNERS', 'FusedMaxIoUAssigner'))
    except Exception:
        if strict:
            raise
    return done
