"""Multi-rank sharding logic (orientedobjectdetection_b200/dist.py) on CPU:
world-size-2 gloo process groups with the CPU oracle injected as the compute
function, checked against the single-process oracle result."""
import copy
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle import ref_api
from orientedobjectdetection_b200 import dist as rdist
from orientedobjectdetection_b200 import synth


def _free_port():
    with socket.socket() as s:
        s.bind(('127.0.0.1', 0))
        return s.getsockname()[1]


def _oracle_overlaps(b1, b2, mode='iou', is_aligned=False):
    return ref_api.rbbox_overlaps(b1, b2, mode, is_aligned)


def _oracle_merge(results, offsets, img_shape, iou_thr=0.1, device='cpu'):
    return ref_api.merge_results(results, offsets, iou_thr)


def _c4_images(n_img, per_level):
    out = []
    for i in range(n_img):
        b, s, ids = synth.c4_rpn_image(seed=i, per_level=per_level + 7 * i, levels=3)
        out.append((torch.from_numpy(b), torch.from_numpy(s), torch.from_numpy(ids)))
    return out


def _worker(rank, world, port, q):
    os.environ['MASTER_ADDR'] = '127.0.0.1'
    os.environ['MASTER_PORT'] = str(port)
    dist.init_process_group('gloo', rank=rank, world_size=world)
    try:
        res = {}
        # --- row-sharded IoU, even and ragged row counts
        for n in (64, 61):
            a, b = synth.c1_iou(seed=3, n=n, m=50, extent=120.0)
            full = rdist.rbbox_overlaps_sharded(torch.from_numpy(a), torch.from_numpy(b),
                                                compute=_oracle_overlaps)
            blk, (lo, hi) = rdist.rbbox_overlaps_sharded(torch.from_numpy(a), torch.from_numpy(b),
                                                         gather=False, compute=_oracle_overlaps)
            res['iou_%d' % n] = full.numpy()
            res['blk_%d' % n] = (blk.numpy(), lo, hi)
        # --- per-image sharded NMS
        imgs = _c4_images(5, 40)
        keeps = rdist.nms_images_sharded(imgs, 0.5, max_per_img=60, keep_fn=ref_api.nms_keep_segmented)
        res['keeps'] = [k.numpy() for k in keeps]
        # --- class-sharded patch merge
        r, offs, shape = synth.c5_patch_results(seed=1, width=2200, height=2200, dets_per_patch=60,
                                                num_classes=5)
        mine = copy.deepcopy(r)
        merged = rdist.merge_results_sharded(mine, offs, shape, 0.1, merge_fn=_oracle_merge)
        res['merged'] = merged
        res['translated'] = mine[1][0]
        q.put((rank, res))
    finally:
        dist.destroy_process_group()


@pytest.fixture(scope='module')
def two_rank_results():
    ctx = mp.get_context('spawn')
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    out = {}
    for _ in procs:
        rank, res = q.get(timeout=240)
        out[rank] = res
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    return out


def test_row_shard_and_lpt():
    assert [rdist.row_shard(10, 4, r) for r in range(4)] == [(0, 3), (3, 6), (6, 8), (8, 10)]
    assert rdist.row_shard(0, 2, 1) == (0, 0)
    owner = rdist.lpt_assign([9, 1, 1, 1, 8, 1], 2)
    loads = [sum(c for c, o in zip([9, 1, 1, 1, 8, 1], owner) if o == r) for r in range(2)]
    assert sorted(loads) == [10, 11]
    assert list(rdist.lpt_assign([5, 5, 5], 1)) == [0, 0, 0]


def test_nms_images_single_process_equals_per_image():
    imgs = _c4_images(3, 50)
    keeps = rdist.nms_images(imgs, 0.5, max_per_img=40, keep_fn=ref_api.nms_keep_segmented)
    for (b, s, ids), k in zip(imgs, keeps):
        _, want = ref_api.batched_nms_rotated(b, s, ids, 0.5)
        assert k.tolist() == want[:40].tolist()
    assert rdist.nms_images([], 0.5) == []


def test_sharded_iou_two_ranks(two_rank_results):
    for n in (64, 61):
        a, b = synth.c1_iou(seed=3, n=n, m=50, extent=120.0)
        want = ref_api.rbbox_overlaps(torch.from_numpy(a), torch.from_numpy(b)).numpy()
        for rank in (0, 1):
            assert np.array_equal(two_rank_results[rank]['iou_%d' % n], want)
            blk, lo, hi = two_rank_results[rank]['blk_%d' % n]
            assert (lo, hi) == rdist.row_shard(n, 2, rank)
            assert np.array_equal(blk, want[lo:hi])


def test_sharded_nms_two_ranks(two_rank_results):
    imgs = _c4_images(5, 40)
    for i, (b, s, ids) in enumerate(imgs):
        _, want = ref_api.batched_nms_rotated(b, s, ids, 0.5)
        for rank in (0, 1):
            assert two_rank_results[rank]['keeps'][i].tolist() == want[:60].tolist()


def test_sharded_merge_two_ranks(two_rank_results):
    r, offs, shape = synth.c5_patch_results(seed=1, width=2200, height=2200, dets_per_patch=60,
                                            num_classes=5)
    mine = copy.deepcopy(r)
    want = ref_api.merge_results(mine, offs, 0.1)
    for rank in (0, 1):
        got = two_rank_results[rank]['merged']
        assert len(got) == len(want)
        for g, w in zip(got, want):
            assert g.shape == w.shape and np.array_equal(g, w)
        # every class's patch arrays were translated in place on every rank
        assert np.array_equal(two_rank_results[rank]['translated'], mine[1][0])
