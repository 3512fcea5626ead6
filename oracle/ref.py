"""ORACLE — TEST INFRASTRUCTURE ONLY (ctypes binding of oracle/librotgeo_ref.so).

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s
``cpu_baseline`` / ``--impl reference`` legs may import this module.  The
product package ``orientedobjectdetection_b200`` never does.

Functions mirror the L1 seam of the reference (``mmcv.ops.box_iou_rotated`` /
``mmcv.ops.nms_rotated`` as called at reference
mmrotate/core/bbox/iou_calculators/rotate_iou2d_calculator.py:89 and
mmrotate/core/post_processing/bbox_nms_rotated.py:80) on numpy arrays.
"""
import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, 'librotgeo_ref.so')
_lib = None


def build(force=False):
    """Compile the C restatement (gcc via oracle/Makefile)."""
    srcs = [os.path.join(_HERE, f) for f in ('rotgeo_ref.c', 'rotgeo_ref_impl.inc', 'Makefile')]
    stale = (not os.path.exists(_SO)) or any(
        os.path.getmtime(s) > os.path.getmtime(_SO) for s in srcs)
    if force or stale:
        subprocess.check_call(['make', '-C', _HERE, '-B', 'librotgeo_ref.so'],
                              stdout=subprocess.DEVNULL)
    return _SO


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_SO):
            build()
        L = ctypes.CDLL(_SO)
        vp, i64, i32, f64 = ctypes.c_void_p, ctypes.c_long, ctypes.c_int, ctypes.c_double
        for sfx, ct in (('f32', ctypes.c_float), ('f64', ctypes.c_double)):
            f = getattr(L, 'rotgeo_ref_iou_' + sfx)
            f.argtypes = [vp, i64, i64, vp, i64, i64, i32, i32, vp, i32]
            f.restype = None
            f = getattr(L, 'rotgeo_ref_nms_' + sfx)
            f.argtypes = [vp, i64, vp, i64, f64, vp]
            f.restype = i64
            f = getattr(L, 'rotgeo_ref_pair_' + sfx)
            f.argtypes = [vp, vp, i32]
            f.restype = ct
        L.rotgeo_ref_max_threads.restype = i32
        _lib = L
    return _lib


def max_threads():
    return int(lib().rotgeo_ref_max_threads())


def _prep(a):
    a = np.asarray(a)
    if a.dtype not in (np.float32, np.float64):
        a = a.astype(np.float32)
    return np.ascontiguousarray(a)


def box_iou_rotated(b1, b2, mode='iou', aligned=False, threads=1):
    """CPU oracle of ``mmcv.ops.box_iou_rotated`` (clockwise flag omitted: it
    negates both angles and cannot change the value).
    Rows may carry extra columns (only the first five are read, as the
    reference relies on at eval_map.py:60-62)."""
    assert mode in ('iou', 'iof')
    b1, b2 = _prep(b1), _prep(b2)
    if b1.dtype != b2.dtype:
        b2 = b2.astype(b1.dtype)
    n1, n2 = b1.shape[0], b2.shape[0]
    if aligned:
        assert n1 == n2
        out = np.zeros((n1,), dtype=b1.dtype)
    else:
        out = np.zeros((n1, n2), dtype=b1.dtype)
    if n1 == 0 or n2 == 0:
        return out
    sfx = 'f32' if b1.dtype == np.float32 else 'f64'
    getattr(lib(), 'rotgeo_ref_iou_' + sfx)(
        b1.ctypes.data, n1, b1.shape[1], b2.ctypes.data, n2, b2.shape[1],
        0 if mode == 'iou' else 1, int(bool(aligned)), out.ctypes.data, int(threads))
    return out


def stable_desc_order(scores):
    """Score-descending order, ties broken by lower index first.  The
    reference sorts with ``scores.sort(0, descending=True)`` (order among ties
    unspecified); both the oracle and the product pin ties this way."""
    scores = np.asarray(scores)
    return np.argsort(-scores.astype(np.float64), kind='stable').astype(np.int64)


def nms_rotated(dets, scores, iou_threshold):
    """CPU oracle of ``mmcv.ops.nms_rotated`` (CPU semantics, ``>=``).
    Returns ``(dets[K,6], keep[K] int64)`` or ``(dets, None)`` when empty."""
    dets = _prep(dets)
    scores = np.ascontiguousarray(np.asarray(scores).astype(dets.dtype))
    n = dets.shape[0]
    if n == 0:
        return dets, None
    order = np.ascontiguousarray(stable_desc_order(scores))
    keep = np.empty((n,), dtype=np.int64)
    sfx = 'f32' if dets.dtype == np.float32 else 'f64'
    k = getattr(lib(), 'rotgeo_ref_nms_' + sfx)(
        dets.ctypes.data, dets.shape[1], order.ctypes.data, n, float(iou_threshold),
        keep.ctypes.data)
    keep = keep[:k].copy()
    out = np.concatenate([dets[keep, :5], scores[keep, None]], axis=1)
    return out, keep


def pair_iou(b1, b2, mode='iou', dtype=np.float32):
    b1 = np.ascontiguousarray(np.asarray(b1, dtype=dtype))
    b2 = np.ascontiguousarray(np.asarray(b2, dtype=dtype))
    sfx = 'f32' if dtype == np.float32 else 'f64'
    return float(getattr(lib(), 'rotgeo_ref_pair_' + sfx)(
        b1.ctypes.data, b2.ctypes.data, 0 if mode == 'iou' else 1))
