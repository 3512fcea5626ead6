"""CPU-side checks of the boundary: the C-ABI library loads and exports every
symbol include/rotgeo.h declares, size queries work without a GPU, and the
python seam keeps the reference's argument/empty-input/error behaviour."""
import ctypes
import os
import re

import numpy as np
import pytest
import torch

import orientedobjectdetection_b200 as rg
from orientedobjectdetection_b200 import _lib, build

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope='module')
def lib():
    build.build()
    return _lib.load()


def test_exports_match_header(lib):
    hdr = open(os.path.join(ROOT, 'include', 'rotgeo.h')).read()
    declared = set(re.findall(r'\b(rotgeo_[a-z_0-9]+)\s*\(', hdr))
    assert declared == set(_lib.SYMBOLS)
    raw = ctypes.CDLL(_lib.SO_PATH)
    for sym in declared:
        assert getattr(raw, sym) is not None
    assert lib.rotgeo_version() == 2000


def test_size_queries_without_gpu(lib):
    assert lib.rotgeo_packed_bytes(0) == 0
    assert lib.rotgeo_packed_bytes(1) == 128 * 96
    assert lib.rotgeo_packed_bytes(129) == 256 * 96
    assert lib.rotgeo_iou_workspace_bytes(8192, 8192) == 2 * 8192 * 96
    assert lib.rotgeo_nms_workspace_bytes(0, None, 0) == 0
    one = lib.rotgeo_nms_workspace_bytes(2000, None, 0)
    offs = np.array([0, 500, 500, 2000], dtype=np.int64)
    seg = lib.rotgeo_nms_workspace_bytes(2000, offs.ctypes.data, 3)
    assert 0 < seg < one            # block-diagonal mask is smaller than the full one


def test_nms_front_end_size_queries_and_bounds(lib):
    """the device-planned NMS is sized on the host: worst-case bounds from the group sizes"""
    sizes = np.zeros(6, dtype=np.int64)
    assert lib.rotgeo_nms_bounds(10000, None, 1, 5, sizes.ctypes.data) == 0
    n_pad, tiles, words, maxb, nseg, n_emit = sizes.tolist()
    C = (10000 + 127) // 128
    assert n_pad == (C + 4) * 128 and maxb == C and nseg == 5 and n_emit == 10000
    assert tiles >= 5 * (16 * 17 // 2) and words >= 5 * 2000 * 2 * 16      # covers 5 levels x 2000
    one = lib.rotgeo_nms_run_bytes(10000, sizes.ctypes.data)
    g = np.array([2000] * 5, dtype=np.int64)
    assert lib.rotgeo_nms_bounds(10000, g.ctypes.data, 5, 1, sizes.ctypes.data) == 0
    assert sizes[0] == 5 * 16 * 128 and sizes[1] == 5 * (16 * 17 // 2) and sizes[4] == 5
    assert 0 < lib.rotgeo_nms_run_bytes(10000, sizes.ctypes.data) < one
    assert lib.rotgeo_nms_sort_bytes(10000, g.ctypes.data, 5, 1) > 3 * 4 * 10000
    assert lib.rotgeo_nms_sort_bytes(10000, g.ctypes.data, 4, 1) == 0        # sizes do not add up
    assert lib.rotgeo_nms_bounds(10000, g.ctypes.data, 4, 1, sizes.ctypes.data) == -1
    # argument errors before any launch
    assert lib.rotgeo_nms_sort(None, 10, 5, None, None, 0, None, 1, 1, None, 0, None) == -1
    ok = np.array([128, 1, 2, 1, 1, 10], dtype=np.int64)
    assert lib.rotgeo_nms_run(None, 10, 5, None, 0, None, 1, 1, None, ok.ctypes.data, 0.0, None, None, None,
                              None, 0, None) == -1
    assert b'positive' in lib.rotgeo_last_error()


def test_argument_errors_do_not_need_a_gpu(lib):
    assert lib.rotgeo_iou_pairwise(None, -1, 5, None, 3, 5, 0, 0.0, None, 3, None, 0, None) == -1
    assert b'iou' in lib.rotgeo_last_error()
    assert lib.rotgeo_iou_pairwise_packed(None, 3, None, 3, 7, None, 3, None) == -1
    assert lib.rotgeo_nms(None, 4, 3, None, None, None, None, 0, 0.1, None, None, None, 0, None) == -1


def test_python_seam_host_logic():
    e5 = torch.zeros((0, 5))
    b = torch.rand(4, 5)
    # empty inputs never reach the device (reference rotate_iou2d_calculator.py:80-81)
    assert rg.rbbox_overlaps(e5, b).shape == (0, 4)
    assert rg.rbbox_overlaps(b, e5).shape == (4, 0)
    assert rg.rbbox_overlaps(e5, e5, is_aligned=True).shape == (0, 1)
    assert rg.box_iou_rotated(e5, b).shape == (0, 4)
    d, k = rg.nms_rotated(e5, torch.zeros(0), 0.1)
    assert k is None and d.shape == (0, 5)
    with pytest.raises(AssertionError):
        rg.rbbox_overlaps(b, b, mode='giou')
    with pytest.raises(AssertionError):
        rg.rbbox_overlaps(torch.rand(4, 4), b)
    with pytest.raises(AssertionError):
        rg.RBboxOverlaps2D()(torch.rand(4, 7), b)
    with pytest.raises(TypeError):
        rg.translate_bboxes(np.zeros((2, 4), np.float32), np.array([1, 2]))
    # multiclass with nothing above the threshold: (0,6) dets, (0,) labels
    dets, labels = rg.multiclass_nms_rotated(torch.rand(6, 5), torch.zeros(6, 4), 0.05,
                                             dict(iou_thr=0.1), 100)
    assert dets.shape == (0, 6) and labels.shape == (0,)
    out, keep = rg.batched_nms(torch.rand(5, 5), torch.rand(5), torch.zeros(5, dtype=torch.long), None)
    assert out.shape == (5, 6) and torch.all(out[:-1, -1] >= out[1:, -1])


@pytest.mark.skipif(torch.cuda.is_available(), reason='checks the no-GPU failure mode')
def test_no_cpu_fallback():
    b = torch.rand(4, 5)
    with pytest.raises(RuntimeError, match='no CPU fallback'):
        rg.box_iou_rotated(b, b)
    with pytest.raises(RuntimeError, match='no CPU fallback'):
        rg.nms_rotated(b, torch.rand(4), 0.1)
