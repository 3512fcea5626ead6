"""The sharding layer on real GPUs over NCCL (run under torchrun, one rank per GPU):
row-sharded IoU with the all-gather of the matrix, images sharded over ranks, classes of a
patch merge sharded over ranks — each compared with the single-GPU result of rank 0's own
library call, plus timings of BASELINE C3 (i) compute, (ii) compute + all-gather, (iii) fused
row/column maxima + all-gather of the maxima.
    python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 tools/dist_check_nccl.py"""
import copy, json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch
import torch.distributed as dist
import orientedobjectdetection_b200 as rg
from orientedobjectdetection_b200 import dist as rdist, synth

rank, world, local = int(os.environ['RANK']), int(os.environ['WORLD_SIZE']), int(os.environ['LOCAL_RANK'])
torch.cuda.set_device(local)
dev = torch.device('cuda', local)
dist.init_process_group('nccl', device_id=dev)
out = {'world': world}
cu = lambda x: torch.from_numpy(x).to(dev)

# ---- IoU rows sharded, even and ragged
for n in (4096, 4001):
    a, b = synth.c3_iou(0, n, 3000, 2048.0)
    full = rdist.rbbox_overlaps_sharded(cu(a), cu(b))
    want = rg.rbbox_overlaps(cu(a), cu(b))
    assert full.shape == want.shape and torch.equal(full, want), n
out['iou_sharded_equal'] = True

# ---- images sharded
imgs = [tuple(cu(x) for x in synth.c4_rpn_image(s, per_level=400 + 30 * s)) for s in range(7)]
got = rdist.nms_images_sharded(imgs, 0.8, 500)
want = rdist.nms_images(imgs, 0.8, 500)
assert all(torch.equal(g.to(dev), w.to(dev)) for g, w in zip(got, want))
out['nms_images_sharded_equal'] = True

# ---- patch merge, classes sharded
res, offs, shape = synth.c5_patch_results(seed=2, width=4000, height=4000, dets_per_patch=200, num_classes=15)
m1 = rdist.merge_results_sharded(copy.deepcopy(res), offs, shape, 0.1)
m0 = rg.merge_results(copy.deepcopy(res), offs, shape, 0.1)
assert all(np.array_equal(x, y) for x, y in zip(m1, m0))
out['merge_results_sharded_equal'] = True

if '--check-only' in sys.argv:
    if rank == 0:
        print(json.dumps(out))
    dist.destroy_process_group()
    sys.exit(0)

# ---- C3 timings: the global (8192 x 8192) matrix row-sharded (strong scaling of ONE matrix)
a, b = synth.c3_iou(0)
ad, bd = cu(a), cu(b)


def timed(fn, reps=10):
    for _ in range(3):
        fn()
    torch.cuda.synchronize(); dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record(); e1.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / reps], device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t)


def maxima_gathered():
    lo, hi = rdist.row_shard(ad.shape[0], world, rank)
    rm, ra, cm, ca = rg.rbbox_overlaps_max(ad[lo:hi], bd)
    rows = torch.empty(ad.shape[0], dtype=torch.float32, device=dev)
    dist.all_gather_into_tensor(rows, rm)
    # column maxima: max over ranks of packed (value, row) keys
    key = (cm.view(torch.int32).to(torch.int64) << 32) | (0xFFFFFFFF - (ca + lo))
    dist.all_reduce(key, op=dist.ReduceOp.MAX)
    return rows, key


out['c3_global_8192x8192_ms'] = {
    'compute_only': timed(lambda: rdist.rbbox_overlaps_sharded(ad, bd, gather=False)),
    'compute_plus_allgather_of_matrix': timed(lambda: rdist.rbbox_overlaps_sharded(ad, bd)),
    'fused_maxima_plus_allgather': timed(maxima_gathered),
}
if rank == 0:
    print(json.dumps(out))
dist.destroy_process_group()
