"""rotgeo_iou_pairwise_static — the entry point that allocates nothing (no counterpart in the
reference): called through the C-ABI and compared
with the existing entry point rotgeo_iou_pairwise on the same inputs, on two calls in a row
(the second on a SMALLER, different problem, so stale contents of the shared static scratch
would show), within the tolerance the IoU parity tests use (1e-5; the kernels are the same, so
the outputs are in fact bit-identical).  Collected last on purpose."""
import numpy as np
import pytest
import torch

from orientedobjectdetection_b200 import _lib, synth

pytestmark = pytest.mark.gpu
IOU_TOL = 1e-5


def _both(L, a, b, mode, min_wh):
    """(static entry point, existing entry point) on device copies of a, b; outputs pre-filled
    with a sentinel so that an element neither call wrote is caught"""
    ad, bd = torch.from_numpy(a).cuda(), torch.from_numpy(b).cuda()
    n1, n2 = a.shape[0], b.shape[0]
    stream = torch.cuda.current_stream().cuda_stream
    got = torch.full((n1, n2), -7.0, dtype=torch.float32, device='cuda')
    want = torch.full((n1, n2), -7.0, dtype=torch.float32, device='cuda')
    _lib.check(L.rotgeo_iou_pairwise_static(ad.data_ptr(), n1, a.shape[1], bd.data_ptr(), n2, b.shape[1], mode,
                                            min_wh, got.data_ptr(), n2, stream), 'rotgeo_iou_pairwise_static')
    nbytes = L.rotgeo_iou_workspace_bytes(n1, n2)
    ws = torch.empty(nbytes, dtype=torch.uint8, device='cuda')
    _lib.check(L.rotgeo_iou_pairwise(ad.data_ptr(), n1, a.shape[1], bd.data_ptr(), n2, b.shape[1], mode, min_wh,
                                     want.data_ptr(), n2, ws.data_ptr(), nbytes, stream), 'rotgeo_iou_pairwise')
    torch.cuda.synchronize()
    return got.cpu().numpy(), want.cpu().numpy()


def test_static_entry_point_matches_workspace_entry_point_on_two_calls_in_a_row():
    L = _lib.load()
    # call 1: 3000 x 2500 boxes, dense scene, IoU, the 1e-3 clamp of rbbox_overlaps
    a, b = synth.c1_iou(seed=11, n=3000, m=2500, extent=700.0)
    got, want = _both(L, a, b, 0, 1e-3)
    assert (want > 0).sum() > 1000 and want.min() >= 0.0
    assert np.abs(got - want).max() <= IOU_TOL
    assert np.array_equal(got, want)
    # call 2, right after: smaller, ragged, other boxes, IoF, 6-column rows — reuses the same scratch
    a2, b2 = synth.c1_iou(seed=12, n=333, m=1001, extent=300.0)
    b2 = np.ascontiguousarray(np.concatenate([b2, np.full((b2.shape[0], 1), 0.5, np.float32)], axis=1))
    got2, want2 = _both(L, a2, b2, 1, 0.0)
    assert (want2 > 0).sum() > 1000 and want2.min() >= 0.0
    assert np.abs(got2 - want2).max() <= IOU_TOL
    assert np.array_equal(got2, want2)
    # and the first problem once more: same answer as before
    got3, _ = _both(L, a, b, 0, 1e-3)
    assert np.array_equal(got3, got)


def test_static_entry_point_limit_and_errors():
    L = _lib.load()
    d = torch.zeros((8, 5), device='cuda')
    o = torch.zeros((8, 8), device='cuda')
    # the stated maximum fits (8192 + 8192 padded boxes) ...
    a, b = synth.c3_iou(seed=3)
    ad, bd = torch.from_numpy(a).cuda(), torch.from_numpy(b).cuda()
    out = torch.empty((8192, 8192), dtype=torch.float32, device='cuda')
    assert L.rotgeo_iou_pairwise_static(ad.data_ptr(), 8192, 5, bd.data_ptr(), 8192, 5, 0, 1e-3, out.data_ptr(),
                                        8192, torch.cuda.current_stream().cuda_stream) == 0
    torch.cuda.synchronize()
    assert float(out.min()) >= 0.0 and float(out.max()) <= 1.0 + 1e-5
    # ... one block more does not: a clear error, nothing launched
    rc = L.rotgeo_iou_pairwise_static(d.data_ptr(), 8193, 5, d.data_ptr(), 8192, 5, 0, 0.0, o.data_ptr(), 8192, None)
    assert rc == -2 and b'static scratch' in L.rotgeo_last_error()
    assert L.rotgeo_iou_pairwise_static(d.data_ptr(), 8, 5, d.data_ptr(), 8, 5, 3, 0.0, o.data_ptr(), 8, None) == -1
    assert L.rotgeo_iou_pairwise_static(d.data_ptr(), 8, 5, d.data_ptr(), 8, 5, 0, 0.0, None, 8, None) == -1
    assert L.rotgeo_iou_pairwise_static(None, 0, 5, None, 5, 5, 0, 0.0, None, 5, None) == 0      # empty: nothing to do
