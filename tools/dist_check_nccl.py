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
# odd column count (scalar push path), IoF, and twice in a row (the window is reused)
a, b = synth.c3_iou(1, 1500, 2999, 900.0)
for _ in range(2):
    full = rdist.rbbox_overlaps_sharded(cu(a), cu(b), 'iof')
    assert torch.equal(full, rg.rbbox_overlaps(cu(a), cu(b), 'iof'))
out['iou_sharded_peer_path'] = bool(rdist.peer.available())
# fused maxima of the row-sharded matrix
for n in (4096, 4001, 3):
    a, b = synth.c3_iou(2, n, 3000, 1024.0)
    got = rdist.rbbox_overlaps_max_sharded(cu(a), cu(b))
    want = rg.rbbox_overlaps_max(cu(a), cu(b))
    assert all(torch.equal(g, w) for g, w in zip(got, want)), n
out['iou_max_sharded_equal'] = True

# ---- images sharded
imgs = [tuple(cu(x) for x in synth.c4_rpn_image(s, per_level=400 + 30 * s)) for s in range(7)]
got = rdist.nms_images_sharded(imgs, 0.8, 500)
want = rdist.nms_images(imgs, 0.8, 500)
assert all(torch.equal(g.to(dev), w.to(dev)) for g, w in zip(got, want))
out['nms_images_sharded_equal'] = True
# every rank holds only ITS images (the others are None), twice in a row (double-buffered window halves)
sizes = [int(b.shape[0]) for b, _, _ in imgs]
owner = rdist.lpt_assign([float(v) ** 2 for v in sizes], world)
local = [im if owner[i] == rank else None for i, im in enumerate(imgs)]
for _ in range(3):
    got = rdist.nms_images_sharded(local, 0.8, 500, sizes=sizes)
    assert all(torch.equal(g.to(dev), w.to(dev)) for g, w in zip(got, want))
out['nms_images_rank_local_equal'] = True
mine_ = [i for i in range(len(imgs)) if owner[i] == rank]


def batched(src, idx, sizes_):
    if not idx:      # a rank without images still takes part in the exchange
        return rdist.Batched(torch.zeros((0, 5), device=dev), torch.zeros(0, device=dev),
                             torch.zeros(0, dtype=torch.int64, device=dev), sizes_)
    return rdist.Batched(torch.cat([src[i][0] for i in idx]), torch.cat([src[i][1] for i in idx]),
                         torch.cat([src[i][2] for i in idx]), sizes_)


bat = batched(imgs, mine_, sizes)
for _ in range(2):
    got = rdist.nms_images_sharded(bat, 0.8, 500)
    assert all(torch.equal(g.to(dev), w.to(dev)) for g, w in zip(got, want))
out['nms_images_batched_equal'] = True

# ---- patch merge, classes sharded
res, offs, shape = synth.c5_patch_results(seed=2, width=4000, height=4000, dets_per_patch=200, num_classes=15)
m1 = rdist.merge_results_sharded(copy.deepcopy(res), offs, shape, 0.1)
m0 = rg.merge_results(copy.deepcopy(res), offs, shape, 0.1)
assert all(np.array_equal(x, y) for x, y in zip(m1, m0))
out['merge_results_sharded_equal'] = True
# device-resident, class-sharded: every rank holds only its classes
shifted = copy.deepcopy(res)
for p_, o_ in zip(shifted, offs):
    for a_ in p_:
        a_[:, :2] += o_
per_cls = [np.concatenate([p_[c] for p_ in shifted], axis=0) for c in range(15)]
n_c = [a_.shape[0] for a_ in per_cls]
own = rdist.lpt_assign([float(v) ** 2 for v in n_c], world)
dets = [cu(per_cls[c]) if own[c] == rank and n_c[c] else None for c in range(15)]
for _ in range(2):
    md = rdist.merge_dets_sharded(dets, n_c, 0.1)
    assert all(np.array_equal(x.cpu().numpy(), y) for x, y in zip(md, m0))
out['merge_dets_sharded_equal'] = True

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
    return rdist.rbbox_overlaps_max_sharded(ad, bd)


out['c3_global_8192x8192_ms'] = {
    'compute_only': timed(lambda: rdist.rbbox_overlaps_sharded(ad, bd, gather=False)),
    'compute_plus_allgather_of_matrix': timed(lambda: rdist.rbbox_overlaps_sharded(ad, bd)),
    'fused_maxima_plus_allgather': timed(maxima_gathered),
}
# ---- C4 / C5 sharded, device-resident, rank-local inputs
def both(fn, reps=5):
    """(CUDA-event ms, wall ms), max over ranks, best of reps"""
    fn()
    ev, wl = [], []
    for _ in range(reps):
        torch.cuda.synchronize(); dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        e0.record(); fn(); e1.record()
        torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1), 1e3 * (time.perf_counter() - t0)], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ev.append(float(t[0])); wl.append(float(t[1]))
    return {'device_ms': min(ev), 'wall_ms': min(wl)}


c4 = [synth.c4_rpn_image(s_) for s_ in range(64)]
sizes = [int(b_.shape[0]) for b_, _, _ in c4]
owner = rdist.lpt_assign([float(v) ** 2 for v in sizes], world)
local = [tuple(cu(x) for x in im) if owner[i] == rank else None for i, im in enumerate(c4)]
out['c4_64img_sharded_rank_local'] = both(lambda: rdist.nms_images_sharded(local, 0.8, 2000, sizes=sizes))
mine_ = [i for i in range(64) if owner[i] == rank]
bat = batched(local, mine_, sizes)
out['c4_64img_sharded_rank_local_batched'] = both(lambda: rdist.nms_images_sharded(bat, 0.8, 2000))
res, offs, shape = synth.c5_patch_results(0)
for p_, o_ in zip(res, offs):
    for a_ in p_:
        a_[:, :2] += o_
per_cls = [np.concatenate([p_[c] for p_ in res], axis=0) for c in range(15)]
n_c = [a_.shape[0] for a_ in per_cls]
own = rdist.lpt_assign([float(v) ** 2 for v in n_c], world)
dets = [cu(per_cls[c]) if own[c] == rank else None for c in range(15)]
out['c5_200k_class_sharded_device_resident'] = both(lambda: rdist.merge_dets_sharded(dets, n_c, 0.1))
if '--profile' in sys.argv and rank == 0:
    from torch.profiler import profile, ProfilerActivity
    for name, fn in (('c4 sharded', lambda: rdist.nms_images_sharded(local, 0.8, 2000, sizes=sizes)),
                     ('c5 sharded', lambda: rdist.merge_dets_sharded(dets, n_c, 0.1))):
        pass
if '--profile' in sys.argv:
    from torch.profiler import profile, ProfilerActivity
    for name, fn in (('c4 sharded', lambda: rdist.nms_images_sharded(bat, 0.8, 2000)),
                     ('c5 sharded', lambda: rdist.merge_dets_sharded(dets, n_c, 0.1))):
        for _ in range(2):
            fn()
        torch.cuda.synchronize(); dist.barrier()
        with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
            for _ in range(5):
                fn()
            torch.cuda.synchronize()
        if rank == 0:
            rows = [(e.key, e.device_time_total / 5, e.cpu_time_total / 5, e.count / 5) for e in prof.key_averages()]
            print('== %s (rank 0): GPU %.1f us per call' % (name, sum(r[1] for r in rows)), file=sys.stderr)
            for k, d, c, n_ in sorted(rows, key=lambda r: -r[1])[:10]:
                print('   gpu %7.1f us  x%.0f  %s' % (d, n_, k[:80]), file=sys.stderr)
            for k, d, c, n_ in sorted(rows, key=lambda r: -r[2])[:12]:
                print('   cpu %7.1f us  x%.0f  %s' % (c, n_, k[:80]), file=sys.stderr)
if '--cprofile' in sys.argv:
    import cProfile, pstats, io
    fn = lambda: rdist.nms_images_sharded(bat, 0.8, 2000)
    for _ in range(3):
        fn()
    torch.cuda.synchronize(); dist.barrier()
    pr = cProfile.Profile()
    pr.enable()
    for _ in range(50):
        fn()
    pr.disable()
    if rank == 0:
        st = io.StringIO()
        pstats.Stats(pr, stream=st).sort_stats('tottime').print_stats(22)
        print(st.getvalue(), file=sys.stderr)
out['c5_class_sizes'] = n_c
out['c5_class_owner'] = [int(v) for v in own]
if rank == 0:
    print(json.dumps(out))
dist.destroy_process_group()
