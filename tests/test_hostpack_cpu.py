"""_hostpack.gather_shift (csrc/hostpack.c) — the C walk over patches x classes of merge_results —
against the numpy walk of the reference (mmrotate/core/patch/merge_results.py:7-29,98-107):
bit-identical rows, identical in-place change of the caller's arrays, refusal of layouts it
does not take (nothing modified then)."""
import numpy as np
import pytest

from orientedobjectdetection_b200 import build

build.build_hostpack()
from orientedobjectdetection_b200 import _hostpack  # noqa: E402


def _scene(seed, patches=7, classes=4, strided=False):
    rng = np.random.default_rng(seed)
    res = []
    for _ in range(patches):
        row = []
        for c in range(classes):
            k = int(rng.integers(0, 9)) if c != 2 else 0          # class 2 is empty everywhere
            a = rng.uniform(-50, 1100, size=(k, 12 if strided else 6)).astype(np.float32)
            row.append(a[:, ::2] if strided else a)
        res.append(row)
    offs = rng.integers(0, 16384, size=(patches, 2)).astype(np.int64)
    return res, offs


def _numpy_walk(res, offs):
    per = []
    for c in range(len(res[0])):
        parts = []
        for i in range(len(res)):
            b = res[i][c]
            b[:, :2] = b[:, :2] + offs[i]                        # the reference's translate_bboxes, (n,6) branch
            parts.append(b)
        per.append(np.concatenate(parts, axis=0))
    return per


@pytest.mark.parametrize('strided', [False, True])
def test_gather_shift_equals_numpy_walk(strided):
    res, offs = _scene(3, strided=strided)
    mine = [[a.copy() if not strided else np.repeat(a, 2, axis=1)[:, ::2] for a in p] for p in res]
    per = _numpy_walk(res, offs)
    out = np.full((1000, 6), -1.0, dtype=np.float32)
    total, counts, filled = _hostpack.gather_shift(mine, offs.astype(np.float64), out)
    assert filled and total == sum(p.shape[0] for p in per) and counts == [p.shape[0] for p in per]
    assert counts[2] == 0
    assert np.array_equal(out[:total], np.concatenate(per, axis=0))
    assert np.all(out[total:] == -1.0)
    for pa, pb in zip(res, mine):                                  # the caller's arrays moved the same way
        for a, b in zip(pa, pb):
            assert np.array_equal(a, b)


def test_gather_shift_float_offsets_and_tiny_values():
    rng = np.random.default_rng(5)
    a = (rng.uniform(-1, 1, size=(64, 6)) * 10.0 ** rng.integers(-30, 4, size=(64, 6))).astype(np.float32)
    offs = np.array([[1234.5, 7.25]])
    want = a.copy()
    want[:, :2] = want[:, :2] + offs[0]
    out = np.empty((64, 6), dtype=np.float32)
    total, counts, filled = _hostpack.gather_shift([[a]], offs, out)
    assert filled and total == 64 and counts == [64]
    assert np.array_equal(out, want) and np.array_equal(a, want)


def test_gather_shift_small_buffer_touches_nothing():
    res, offs = _scene(4)
    before = [[a.copy() for a in p] for p in res]
    out = np.zeros((2, 6), dtype=np.float32)
    total, counts, filled = _hostpack.gather_shift(res, offs.astype(np.float64), out)
    assert not filled and total == sum(counts) > 2
    assert np.all(out == 0)
    assert all(np.array_equal(a, b) for pa, pb in zip(res, before) for a, b in zip(pa, pb))


def test_gather_shift_refuses_other_layouts():
    out = np.zeros((100, 6), dtype=np.float32)
    offs = np.zeros((1, 2))
    good = np.ones((3, 6), dtype=np.float32)
    ro = np.ones((3, 6), dtype=np.float32)
    ro.setflags(write=False)
    for bad in (np.ones((3, 5), dtype=np.float32), np.ones((3, 6), dtype=np.float64), np.ones(6, dtype=np.float32), ro,
                [[1.0] * 6]):
        g = good.copy()
        assert _hostpack.gather_shift([[g, bad]], offs + 5.0, out) is None
        assert np.all(g == 1.0)                                    # validated before anything is shifted
    with pytest.raises(ValueError):
        _hostpack.gather_shift([[good]], np.zeros((2, 2)), out)
    with pytest.raises(ValueError):
        _hostpack.gather_shift([[good], [good, good]], np.zeros((2, 2)), out)
    assert _hostpack.gather_shift([], np.zeros((0, 2)), out) == (0, [], True)


def test_bind_host_to_gpu_is_harmless_without_a_gpu():
    """peer.bind_host_to_gpu(): returns the CPU list it pinned the thread to, or None when NVML / CUDA is not
    there (this container) — never raises, never leaves the affinity empty"""
    import os

    from orientedobjectdetection_b200 import peer
    before = os.sched_getaffinity(0)
    got = peer.bind_host_to_gpu(0)
    assert got is None or (isinstance(got, list) and len(got) > 0)
    assert len(os.sched_getaffinity(0)) > 0
    if got is None:
        assert os.sched_getaffinity(0) == before


def test_gather_shift_large_scene_threads_and_aliased_arrays():
    """above 20 000 rows the copy pass runs on a few threads: same bytes as the numpy walk; an array that appears
    twice is refused (the numpy walk reproduces the reference's double shift)"""
    rng = np.random.default_rng(9)
    res = [[rng.uniform(0, 900, size=(int(rng.integers(20, 60)), 6)).astype(np.float32) for _ in range(5)] for _ in range(150)]
    offs = rng.integers(0, 16000, size=(150, 2)).astype(np.int64)
    mine = [[a.copy() for a in p] for p in res]
    per = _numpy_walk(res, offs)
    out = np.empty((sum(p.shape[0] for p in per) + 7, 6), dtype=np.float32)
    total, counts, filled = _hostpack.gather_shift(mine, offs.astype(np.float64), out)
    assert filled and total >= 20000 and counts == [p.shape[0] for p in per]
    assert np.array_equal(out[:total], np.concatenate(per, axis=0))
    assert all(np.array_equal(a, b) for pa, pb in zip(res, mine) for a, b in zip(pa, pb))
    # aliased: patch 1 reuses patch 0's array object for class 0 -> refused untouched (the reference collects references,
    # shifts the array twice and concatenates its final state twice: the numpy walk of merge_results does that)
    mine[1][0] = mine[0][0]
    before = mine[0][0].copy()
    assert _hostpack.gather_shift(mine, offs.astype(np.float64), out) is None
    assert np.array_equal(mine[0][0], before)
