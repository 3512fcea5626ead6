"""ctypes loader of librotgeo.so (the C-ABI declared in include/rotgeo.h).

There is no fallback of any kind: if the library is missing or a call fails the
caller gets a RuntimeError.  The CPU oracle under oracle/ is never imported here.
"""
import ctypes
import os

_PKG = os.path.dirname(os.path.abspath(__file__))
SO_PATH = os.path.join(_PKG, 'librotgeo.so')

# every symbol include/rotgeo.h declares (tests check the export list against this)
SYMBOLS = (
    'rotgeo_version', 'rotgeo_last_error', 'rotgeo_packed_bytes', 'rotgeo_pack_boxes',
    'rotgeo_iou_workspace_bytes', 'rotgeo_iou_pairwise', 'rotgeo_iou_pairwise_packed',
    'rotgeo_iou_aligned', 'rotgeo_iou_pairwise_static', 'rotgeo_nms_workspace_bytes', 'rotgeo_nms',
    'rotgeo_iou_max_workspace_bytes', 'rotgeo_iou_rowcol_max', 'rotgeo_max_iou_assign',
    'rotgeo_eval_tpfp_workspace_bytes', 'rotgeo_eval_tpfp',
    'rotgeo_diff_iou_aligned',
    'rotgeo_nms_sort_bytes', 'rotgeo_nms_sort', 'rotgeo_nms_bounds', 'rotgeo_nms_read_plan',
    'rotgeo_nms_run_bytes', 'rotgeo_nms_run', 'rotgeo_nms_grouped',
    'rotgeo_multiclass_nms_workspace_bytes', 'rotgeo_multiclass_nms',
    'rotgeo_peer_pad_bytes', 'rotgeo_peer_barrier', 'rotgeo_dist_iou_allgather', 'rotgeo_dist_rowcol_stage_bytes',
    'rotgeo_dist_iou_rowcol_max', 'rotgeo_dist_keeps_bytes', 'rotgeo_dist_push_keeps',
)
ERR_RANGE = -4
ERR_WORKSPACE = -2

_lib = None


class RotgeoError(RuntimeError):
    pass


def load():
    """dlopen librotgeo.so and declare the prototypes.  Raises if it is not built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(SO_PATH):
        raise RotgeoError(
            'librotgeo.so is not built (%s). Run `python -m orientedobjectdetection_b200.build` '
            '(needs nvcc). There is no CPU fallback.' % SO_PATH)
    L = ctypes.CDLL(SO_PATH)
    vp, i64, i32, f32, sz = (ctypes.c_void_p, ctypes.c_int64, ctypes.c_int, ctypes.c_float,
                             ctypes.c_size_t)
    L.rotgeo_version.restype = i32
    L.rotgeo_last_error.restype = ctypes.c_char_p
    L.rotgeo_packed_bytes.argtypes = [i64]
    L.rotgeo_packed_bytes.restype = sz
    L.rotgeo_pack_boxes.argtypes = [vp, i64, i32, vp, vp, f32, f32, vp, vp]
    L.rotgeo_pack_boxes.restype = i32
    L.rotgeo_iou_workspace_bytes.argtypes = [i64, i64]
    L.rotgeo_iou_workspace_bytes.restype = sz
    L.rotgeo_iou_pairwise.argtypes = [vp, i64, i32, vp, i64, i32, i32, f32, vp, i64, vp, sz, vp]
    L.rotgeo_iou_pairwise.restype = i32
    L.rotgeo_iou_pairwise_packed.argtypes = [vp, i64, vp, i64, i32, vp, i64, vp]
    L.rotgeo_iou_pairwise_packed.restype = i32
    L.rotgeo_iou_pairwise_static.argtypes = [vp, i64, i32, vp, i64, i32, i32, f32, vp, i64, vp]
    L.rotgeo_iou_pairwise_static.restype = i32
    L.rotgeo_iou_aligned.argtypes = [vp, vp, i64, i32, i32, i32, f32, vp, vp]
    L.rotgeo_iou_aligned.restype = i32
    L.rotgeo_nms_workspace_bytes.argtypes = [i64, vp, i64]
    L.rotgeo_nms_workspace_bytes.restype = sz
    L.rotgeo_nms.argtypes = [vp, i64, i32, vp, vp, vp, vp, i64, f32, vp, vp, vp, sz, vp]
    L.rotgeo_nms.restype = i32
    L.rotgeo_iou_max_workspace_bytes.argtypes = [i64, i64]
    L.rotgeo_iou_max_workspace_bytes.restype = sz
    L.rotgeo_iou_rowcol_max.argtypes = [vp, i64, i32, vp, i64, i32, i32, f32, vp, vp, vp, vp, vp, sz, vp]
    L.rotgeo_iou_rowcol_max.restype = i32
    L.rotgeo_max_iou_assign.argtypes = [vp, i64, i32, vp, i64, i32, i32, f32, f32, f32, f32, f32, i32, i32,
                                        vp, vp, vp, sz, vp]
    L.rotgeo_max_iou_assign.restype = i32
    L.rotgeo_eval_tpfp_workspace_bytes.argtypes = [i64, i64]
    L.rotgeo_eval_tpfp_workspace_bytes.restype = sz
    L.rotgeo_eval_tpfp.argtypes = [vp, i64, i32, vp, vp, i64, i32, vp, vp, i64, f32, vp, vp, vp, vp, sz, vp]
    L.rotgeo_eval_tpfp.restype = i32
    L.rotgeo_diff_iou_aligned.argtypes = [vp, vp, i64, i32, i32, vp, vp, vp, vp]
    L.rotgeo_diff_iou_aligned.restype = i32
    L.rotgeo_nms_sort_bytes.argtypes = [i64, vp, i64, i64]
    L.rotgeo_nms_sort_bytes.restype = sz
    L.rotgeo_nms_sort.argtypes = [vp, i64, i32, vp, vp, i32, vp, i64, i64, vp, sz, vp]
    L.rotgeo_nms_sort.restype = i32
    L.rotgeo_nms_bounds.argtypes = [i64, vp, i64, i64, vp]
    L.rotgeo_nms_bounds.restype = i32
    L.rotgeo_nms_read_plan.argtypes = [vp, i64, vp, i64, i64, vp, vp]
    L.rotgeo_nms_read_plan.restype = i32
    L.rotgeo_nms_run_bytes.argtypes = [i64, vp]
    L.rotgeo_nms_run_bytes.restype = sz
    L.rotgeo_nms_run.argtypes = [vp, i64, i32, vp, i32, vp, i64, i64, vp, vp, f32, vp, vp, vp, vp, sz, vp]
    L.rotgeo_nms_run.restype = i32
    L.rotgeo_nms_grouped.argtypes = [vp, i64, i32, vp, vp, i32, vp, vp, i64, i64, f32, vp, vp, i32, vp, vp, sz, vp, sz, vp, vp]
    L.rotgeo_nms_grouped.restype = i32
    L.rotgeo_multiclass_nms_workspace_bytes.argtypes = [i64, i32]
    L.rotgeo_multiclass_nms_workspace_bytes.restype = sz
    L.rotgeo_multiclass_nms.argtypes = [vp, i32, vp, vp, i64, i32, f32, f32, i64, vp, vp, vp, vp, vp, sz, vp]
    L.rotgeo_multiclass_nms.restype = i32
    u64 = ctypes.c_uint64
    L.rotgeo_peer_pad_bytes.restype = sz
    L.rotgeo_peer_barrier.argtypes = [vp, i32, i32, u64, vp]
    L.rotgeo_peer_barrier.restype = i32
    L.rotgeo_dist_iou_allgather.argtypes = [vp, i32, i32, u64, sz, vp, i64, i64, i64, i32, vp, i64, i32, i32, f32,
                                            vp, sz, vp]
    L.rotgeo_dist_iou_allgather.restype = i32
    L.rotgeo_dist_rowcol_stage_bytes.argtypes = [i64, i64, i32]
    L.rotgeo_dist_rowcol_stage_bytes.restype = sz
    L.rotgeo_dist_iou_rowcol_max.argtypes = [vp, i32, i32, u64, sz, vp, i64, i64, i64, i32, vp, i64, i32, i32, f32,
                                             vp, vp, vp, vp, vp, sz, vp]
    L.rotgeo_dist_iou_rowcol_max.restype = i32
    L.rotgeo_dist_keeps_bytes.argtypes = [i64, i64, i32, i32]
    L.rotgeo_dist_keeps_bytes.restype = sz
    L.rotgeo_dist_push_keeps.argtypes = [vp, i32, i32, u64, sz, vp, vp, i32, vp, i64, i64, vp, i32, i32, vp, vp]
    L.rotgeo_dist_push_keeps.restype = i32
    _lib = L
    return L


def check(rc, what):
    if rc != 0:
        msg = load().rotgeo_last_error().decode('utf-8', 'replace')
        raise RotgeoError('%s failed (code %d): %s' % (what, rc, msg))
