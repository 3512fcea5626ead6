"""Multi-GPU sharding of the rotated-geometry hot path:
one process per GPU, ``torch.distributed`` (NCCL over NVLink on the box, gloo
in the CPU tests) used ONLY to gather results — every unit of work (an IoU
row block, one image's NMS, one class of a patch merge) is independent, so no
collective sits inside a kernel's critical path.

    * rbbox_overlaps_sharded   rows [r*N/G, (r+1)*N/G) x all columns on rank r,
                               optional all-gather of the row blocks
    * nms_images_sharded       image i -> rank by LPT on its box count; each rank
                               runs ALL its images x levels as one segmented
                               launch; padded all-gather of the keep lists
    * merge_results_sharded    class c -> rank by LPT on n_c^2 (pair count);
                               all-gather of the per-class kept rows

The reference has no multi-GPU version of this path: it loops over images in
python inside each DDP rank (rotated_anchor_head.py:581-602,
rotated_rpn_head.py:414-426) and over classes in merge_results
(merge_results.py:90).

``compute`` arguments exist so the CPU tests can drive the sharding logic with
the oracle under gloo; the default is always the CUDA path.
"""
import numpy as np
import torch
import torch.distributed as dist

from . import core, ops


# ------------------------------------------------------------------ partitioning
def row_shard(n, world, rank):
    """contiguous balanced row range of rank ``rank``: sizes differ by at most 1"""
    base, rem = divmod(int(n), int(world))
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def lpt_assign(costs, world):
    """Longest-processing-time-first assignment of units to ranks.  Returns an
    int array ``owner[i]``; deterministic (ties: lower unit index first, lower
    rank first), so every rank computes the same plan without communication."""
    costs = np.asarray(costs, dtype=np.float64)
    owner = np.zeros(costs.shape[0], dtype=np.int64)
    load = np.zeros(world, dtype=np.float64)
    for i in np.argsort(-costs, kind='stable'):
        r = int(np.argmin(load))
        owner[i] = r
        load[r] += costs[i]
    return owner


def _world(group):
    if dist.is_available() and dist.is_initialized():
        return dist.get_world_size(group), dist.get_rank(group)
    return 1, 0


def _comm_device(group, like=None):
    """tensors handed to collectives must live where the backend wants them"""
    if dist.is_initialized() and dist.get_backend(group) == 'nccl':
        return torch.device('cuda', torch.cuda.current_device())
    return torch.device('cpu')


def _all_gather_padded(t, count_max, group):
    """all-gather of 1-D/2-D tensors with a ragged first dim.  Every rank pads
    to ``count_max`` rows; returns the list of un-padded per-rank tensors."""
    world, _ = _world(group)
    n = t.shape[0]
    pad = torch.zeros((count_max,) + tuple(t.shape[1:]), dtype=t.dtype, device=t.device)
    pad[:n] = t
    counts = torch.zeros(world, dtype=torch.int64, device=t.device)
    mine = torch.tensor([n], dtype=torch.int64, device=t.device)
    dist.all_gather_into_tensor(counts, mine, group=group)
    out = torch.empty((world * count_max,) + tuple(t.shape[1:]), dtype=t.dtype, device=t.device)
    dist.all_gather_into_tensor(out, pad, group=group)
    counts = counts.tolist()
    return [out[r * count_max:r * count_max + counts[r]] for r in range(world)]


# ------------------------------------------------------------------ IoU, row-sharded
def rbbox_overlaps_sharded(bboxes1, bboxes2, mode='iou', gather=True, group=None, compute=None):
    """Row-sharded ``rbbox_overlaps`` (BASELINE C3).  Every rank holds the same
    ``bboxes1`` (N,5) and ``bboxes2`` (M,5); rank r computes rows
    ``row_shard(N, G, r)`` against all columns.

    gather=True  -> the full (N, M) matrix on every rank (one all-gather of
                    N/G x M blocks: 268 MB total at 8192^2);
    gather=False -> ``(block, (lo, hi))``, nothing is communicated.
    """
    compute = compute or core.rbbox_overlaps
    world, rank = _world(group)
    n, m = bboxes1.size(0), bboxes2.size(0)
    lo, hi = row_shard(n, world, rank)
    block = compute(bboxes1[lo:hi], bboxes2, mode, False)
    if not gather:
        return block, (lo, hi)
    if world == 1:
        return block
    cdev = _comm_device(group)
    block = block.to(cdev)
    rows_max = row_shard(n, world, 0)[1]            # rank 0 has the largest share
    if n % world == 0:
        out = torch.empty((n, m), dtype=block.dtype, device=cdev)
        dist.all_gather_into_tensor(out, block.contiguous(), group=group)
    else:
        out = torch.cat(_all_gather_padded(block, rows_max, group), dim=0)
    return out.to(bboxes1.device) if out.device != bboxes1.device and not bboxes1.is_cuda else out


# ------------------------------------------------------------------ many-image NMS
def nms_images(images, iou_threshold, max_per_img=-1, keep_fn=None):
    """Class/level-aware rotated NMS of SEVERAL images in one segmented launch.

    ``images``: list of ``(boxes (n_i,5), scores (n_i,), idxs (n_i,))`` — what
    the reference feeds ``batched_nms`` one image at a time
    (oriented_rpn_head.py:275, rotated_rpn_head.py:515).  Returns one int64
    keep tensor per image (indices into that image's boxes, score-descending,
    cut to ``max_per_img``), identical to calling ``ops.batched_nms`` per image.
    """
    if not images:
        return []
    sizes = [int(b.shape[0]) for b, _, _ in images]
    if sum(sizes) == 0:
        return [torch.zeros(0, dtype=torch.int64) for _ in images]
    dev = images[0][0].device
    boxes = torch.cat([b for b, _, _ in images], dim=0)
    scores = torch.cat([s for _, s, _ in images], dim=0)
    ids = torch.cat([i.to(torch.int64) for _, _, i in images], dim=0)
    if keep_fn is None:
        # the product path: offsets, orders, NMS and the split into images all on the device
        # (csrc/rotgeo_frontend.cu "batch of images"); None = ids too spread out for it
        got = ops._nms_keep_images(boxes, scores, ids, sizes, iou_threshold)
        if got is not None:
            keep, counts = got
            out = list(torch.split(keep.to(dev), counts))
            return [k[:max_per_img] for k in out] if max_per_img > 0 else out
        keep_fn = ops._nms_keep
    nimg = len(images)
    total = sum(sizes)
    sizes_t = torch.tensor(sizes, dtype=torch.int64)
    sizes_d = sizes_t.to(dev)
    img_of = torch.repeat_interleave(torch.arange(nimg, dtype=torch.int64, device=dev), sizes_d,
                                     output_size=total)
    # per-image offset trick, same float arithmetic as mmcv.ops.batched_nms on every image:
    # id * (max(cx,cy) + max(w,h) + 1) with the maxima of THAT image (images are contiguous
    # row ranges: one segmented max per term; empty images never index the result)
    m_xy = torch.segment_reduce(boxes[:, :2].max(dim=1).values, 'max', lengths=sizes_d, unsafe=True)
    m_wh = torch.segment_reduce(boxes[:, 2:4].max(dim=1).values, 'max', lengths=sizes_d, unsafe=True)
    offsets = ids.to(boxes) * ((m_xy + m_wh)[img_of] + torch.tensor(1).to(boxes))
    # one segment per (image, id)
    id_lo = ids.min()
    seg = img_of * (ids.max() - id_lo + 1) + (ids - id_lo)
    keep = keep_fn(boxes, scores, iou_threshold, seg_ids=seg, offsets=offsets).to(dev)
    # global score order -> per image (stable: score order survives inside an image)
    kimg = img_of[keep]
    perm = torch.sort(kimg, stable=True).indices
    keep, kimg = keep[perm], kimg[perm]
    starts = (torch.cumsum(sizes_t, 0) - sizes_t).to(dev)
    local = keep - starts[kimg]
    counts = torch.bincount(kimg, minlength=nimg).tolist()
    out = list(torch.split(local, counts))
    return [k[:max_per_img] for k in out] if max_per_img > 0 else out


def nms_images_sharded(images, iou_threshold, max_per_img=-1, group=None, keep_fn=None):
    """``nms_images`` with the images spread over the ranks (BASELINE C4).
    Every rank passes the SAME list; image i is processed by
    ``lpt_assign([n_i^2], G)[i]`` and every rank returns all keep lists."""
    world, rank = _world(group)
    if world == 1:
        return nms_images(images, iou_threshold, max_per_img, keep_fn)
    owner = lpt_assign([float(b.shape[0]) ** 2 for b, _, _ in images], world)
    mine = [i for i in range(len(images)) if owner[i] == rank]
    keeps = nms_images([images[i] for i in mine], iou_threshold, max_per_img, keep_fn)
    cdev = _comm_device(group)
    # ONE collective and ONE device->host read: every rank sends a fixed-size int64 message
    # [kept count of each of its images (H slots) | kept indices, image after image]; the
    # headers of all ranks come to the host in one copy, the index lists stay on the device
    # and are handed out as views.
    idx_of = [[i for i in range(len(images)) if owner[i] == r] for r in range(world)]
    H = max(max(len(v) for v in idx_of), 1)
    cap = 1
    for r in range(world):
        per = [min(int(images[i][0].shape[0]), max_per_img) if max_per_img > 0
               else int(images[i][0].shape[0]) for i in idx_of[r]]
        cap = max(cap, sum(per))
    msg = torch.zeros(H + cap, dtype=torch.int64, device=cdev)
    counts = [int(k.numel()) for k in keeps]
    if counts:
        msg[:len(counts)] = torch.tensor(counts, dtype=torch.int64)
        total = sum(counts)
        if total:
            msg[H:H + total] = torch.cat([k.to(cdev) for k in keeps])
    out_all = torch.empty(world * (H + cap), dtype=torch.int64, device=cdev)
    dist.all_gather_into_tensor(out_all, msg, group=group)
    out_all = out_all.view(world, H + cap)
    heads = out_all[:, :H].cpu().tolist()
    out = [None] * len(images)
    for r in range(world):
        nr = len(idx_of[r])
        if nr == 0:
            continue
        cr = heads[r][:nr]
        parts = torch.split(out_all[r, H:H + sum(cr)], cr)
        for i, part in zip(idx_of[r], parts):
            out[i] = part.to(images[i][0].device)
    return out


# ------------------------------------------------------------------ huge-image merge
def merge_results_sharded(results, offsets, img_shape, iou_thr=0.1, device='cpu', group=None,
                          merge_fn=None):
    """``core.merge_results`` with the classes spread over the ranks (BASELINE
    C5).  Every rank passes the same ``results``; class c is merged by rank
    ``lpt_assign([n_c^2], G)[c]`` and the kept rows are all-gathered, so every
    rank returns the complete list over classes.  Like the reference
    (merge_results.py:22-25) the caller's arrays are shifted in place."""
    merge_fn = merge_fn or core.merge_results
    world, rank = _world(group)
    if world == 1:
        return merge_fn(results, offsets, img_shape, iou_thr, device)
    num_patches, num_classes = len(results), len(results[0])
    n_c = [sum(results[p][c].shape[0] for p in range(num_patches)) for c in range(num_classes)]
    owner = lpt_assign([float(v) ** 2 for v in n_c], world)
    mine = [c for c in range(num_classes) if owner[c] == rank]
    # this rank's classes only; translation (in place) happens for ALL classes so the
    # observable side effect on the caller's arrays matches the single-process call
    sub = [[results[p][c] for c in mine] for p in range(num_patches)]
    merged_mine = merge_fn(sub, offsets, img_shape, iou_thr, device) if mine else []
    for c in range(num_classes):
        if owner[c] != rank:
            for p in range(num_patches):
                core.translate_bboxes(results[p][c], offsets[p])
    cdev = _comm_device(group)
    width = max((results[0][c].shape[1] for c in range(num_classes)), default=6)
    rows = [torch.from_numpy(np.ascontiguousarray(m, dtype=np.float32)).reshape(-1, width)
            for m in merged_mine]
    flat = torch.cat(rows).to(cdev) if rows else torch.zeros((0, width), dtype=torch.float32, device=cdev)
    cnt = torch.tensor([m.shape[0] for m in merged_mine], dtype=torch.int64, device=cdev)
    cap_rows = max(max(sum(n_c[c] for c in range(num_classes) if owner[c] == r) for r in range(world)), 1)
    g_rows = _all_gather_padded(flat, cap_rows, group)
    g_cnt = _all_gather_padded(cnt, max(num_classes, 1), group)
    out = [None] * num_classes
    for r in range(world):
        pos = 0
        cls_r = [c for c in range(num_classes) if owner[c] == r]
        for j, c in enumerate(cls_r):
            k = int(g_cnt[r][j])
            out[c] = g_rows[r][pos:pos + k].cpu().numpy().astype(results[0][c].dtype, copy=False)
            pos += k
    return out
