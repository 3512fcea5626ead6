#!/usr/bin/env python
"""bench.py — rotated IoU Gpairs/s (headline, BASELINE.json C3: 8192x8192 le90
boxes, one such block per GPU) and rotated-NMS boxes/s (secondary, C2/C4), next
to the CPU restatement of the reference's mmcv path timed on this box's cores.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl rotgeo|reference]

One JSON line on stdout (rank 0).  See DESIGN.md "Measurement" for definitions.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

N_ROWS = 8192          # rows per GPU (weak scaling: the global matrix is (N_ROWS*G) x N_COLS)
N_COLS = 8192
EXTENT = 4096.0
WORKLOAD = ('C3 rbbox_overlaps 8192x8192 le90 boxes, extent 4096, synth.c3_iou(seed 0), w,h clamp 1e-3: one such '
            'block per GPU (weak scaling: rank r > 0 draws its rows with seed 1000 + r)')
FLOP_PER_PAIR = 450.0  # mandatory part of the reference algorithm
FP32_PEAK_TFLOPS = 148 * 128 * 2 * 1.965e9 / 1e12   # 74.5, theoretical at max boost


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=20)
    ap.add_argument('--warmup', type=int, default=5)
    ap.add_argument('--impl', default='rotgeo', choices=['rotgeo', 'reference'])
    ap.add_argument('--no-secondary', action='store_true')
    ap.add_argument('--no-cpu-baseline', action='store_true')
    return ap.parse_args()


def peaks():
    p = os.path.join(ROOT, 'MEASURED_PEAKS.json')
    if os.path.exists(p):
        d = json.load(open(p))
        return float(d['hbm_gbs']), 'measured (MEASURED_PEAKS.json hbm_gbs, burst copy)'
    return 6650.0, 'fallback (B200_PROFILING.md)'


class ClockSampler(object):
    """nvidia-smi clocks / throttle reasons sampled while the timed region runs"""
    Q = ('clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,'
         'clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,'
         'clocks_event_reasons.sw_power_cap')

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ['nvidia-smi', '-i', str(self.index), '--query-gpu=' + self.Q,
                 '--format=csv,noheader,nounits', '-lms', '100'],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(',')])

    def stop(self):
        if self.proc is None:
            return {'sm_mhz': None, 'sm_max_mhz': None, 'reasons': ['nvidia-smi unavailable']}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ['hw_slowdown', 'hw_thermal_slowdown', 'sw_thermal_slowdown', 'sw_power_cap']
        for r in self.rows:
            try:
                sm.append(float(r[0]))
                mx.append(float(r[1]))
            except Exception:
                continue
            for nm, v in zip(names, r[3:7]):
                if v.lower().startswith('active'):
                    reasons.add(nm)
        return {'sm_mhz': float(np.median(sm)) if sm else None,
                'sm_max_mhz': float(max(mx)) if mx else None,
                'samples': len(sm), 'reasons': sorted(reasons)}


def cpu_iou_sample(a, b, rows, threads, budget_s):
    """oracle on a row slice of the same workload, repeated until ~budget_s of CPU time is
    spent; returns (best Gpairs/s, seconds spent, repetitions)"""
    from oracle import ref
    ref.box_iou_rotated(a[:64], b, threads=threads)       # warm-up
    best, spent, reps = 0.0, 0.0, 0
    while spent < budget_s and reps < 64:
        t0 = time.perf_counter()
        ref.box_iou_rotated(a[:rows], b, threads=threads)
        dt = time.perf_counter() - t0
        best = max(best, rows * b.shape[0] / dt / 1e9)
        spent += dt
        reps += 1
    return best, spent, reps


def run_reference(args, rank):
    """--impl reference: the CPU restatement of the reference's mmcv path (the
    real mmcv cannot be installed here), all host threads, rank 0 only."""
    if rank != 0:
        return
    from oracle import ref
    from orientedobjectdetection_b200 import synth
    a, b = synth.c3_iou(0, N_ROWS, N_COLS, EXTENT)
    a[:, 2:4] = np.maximum(a[:, 2:4], 1e-3)
    b[:, 2:4] = np.maximum(b[:, 2:4], 1e-3)
    threads = min(ref.max_threads(), len(os.sched_getaffinity(0)))   # not OMP_NUM_THREADS (torchrun sets 1)
    rows = N_ROWS if threads >= 4 else 2048      # a step stays below ~2 s of CPU time
    for _ in range(max(1, min(args.warmup, 2))):
        ref.box_iou_rotated(a[:rows], b, threads=threads)
    times = []
    for s in range(args.steps):
        lo = (s * rows) % N_ROWS
        t0 = time.perf_counter()
        ref.box_iou_rotated(a[lo:lo + rows], b, threads=threads)
        times.append(time.perf_counter() - t0)
    ms = 1e3 * float(np.mean(times))
    val = rows * N_COLS / (ms * 1e-3) / 1e9
    sample = '%d-row slice x %d columns of the 8192x8192 C3 matrix per step' % (rows, N_COLS)
    emit({
        'impl': 'reference', 'metric': 'rotated_iou_gpairs_per_s', 'value': val, 'unit': 'Gpairs/s',
        'n_gpus': args.gpus, 'steps': args.steps, 'warmup': args.warmup, 'ms_per_step': ms,
        'higher_is_better': True, 'scaling': 'weak', 'collective': 'none', 'vs_baseline': None, 'dtype': 'f32',
        'data': 'synthetic',
        'config': {'workload': WORKLOAD, 'sample': sample},
        'cpu_baseline': {'value': val, 'unit': 'Gpairs/s', 'cores': threads, 'kind': 'port',
                         'sample': sample},
        'e2e': {'value': val, 'unit': 'Gpairs/s', 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0},
    })


def timed_steps(torch, fn, steps, flush, stream_sync):
    """per-step CUDA events on the current stream; L2 is flushed between steps
    OUTSIDE the events.  returns list of ms."""
    out = []
    for _ in range(steps):
        flush()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        e1.synchronize()
        out.append(e0.elapsed_time(e1))
    return out


def secondary_nms(torch, rg, synth, reps=5):
    """rotated-NMS boxes/s on one GPU through the public python API, wall clock incl. the
    call's own device->host reads.  C2/C4: inputs resident on the device (as the detector
    heads have them); C5: host numpy lists in, numpy out (as huge_image_demo.py has them)."""
    from orientedobjectdetection_b200 import dist as rdist
    res = {}

    def best(fn, reps=reps):
        fn()
        ts = []
        for _ in range(reps):
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            fn()
            torch.cuda.synchronize()
            ts.append(time.perf_counter() - t0)
        return min(ts)

    for name, dense in (('c2_multiclass_nms_2000x15_sparse', False), ('c2_multiclass_nms_2000x15_dense', True)):
        mb, ms = synth.c2_multiclass(0, dense=dense)
        mbd, msd = torch.from_numpy(mb).cuda(), torch.from_numpy(ms).cuda()
        ncand = int((ms[:, :-1] > 0.05).sum())
        t = best(lambda: rg.multiclass_nms_rotated(mbd, msd, 0.05, dict(iou_thr=0.1), 2000))
        res[name] = {'boxes': ncand, 'ms': 1e3 * t, 'boxes_per_s': ncand / t}
    imgs = [tuple(torch.from_numpy(x).cuda() for x in synth.c4_rpn_image(s)) for s in range(64)]
    bd, sd, idd = imgs[0]
    cfg = dict(type='nms_rotated', iou_threshold=0.8)
    t = best(lambda: rg.batched_nms(bd, sd, idd, cfg))
    res['c4_batched_nms_1img_5x2000'] = {'boxes': int(bd.shape[0]), 'ms': 1e3 * t, 'boxes_per_s': bd.shape[0] / t}
    nb = sum(int(b.shape[0]) for b, _, _ in imgs)
    t = best(lambda: [rg.batched_nms(b, s, i, cfg)[1][:2000] for b, s, i in imgs], 2)
    res['c4_64img_per_image_calls'] = {'boxes': nb, 'ms': 1e3 * t, 'boxes_per_s': nb / t}
    t = best(lambda: rdist.nms_images(imgs, 0.8, 2000), 3)
    res['c4_64img_one_segmented_launch'] = {'boxes': nb, 'ms': 1e3 * t, 'boxes_per_s': nb / t}
    bat = rdist.Batched(torch.cat([b for b, _, _ in imgs]), torch.cat([s for _, s, _ in imgs]),
                        torch.cat([i for _, _, i in imgs]), [int(b.shape[0]) for b, _, _ in imgs])
    t = best(lambda: rdist.nms_images(bat, 0.8, 2000), 3)
    res['c4_64img_one_segmented_launch_batched_input'] = {'boxes': nb, 'ms': 1e3 * t, 'boxes_per_s': nb / t}
    del bat
    del imgs
    n = 30000
    b1 = torch.from_numpy(synth.le90_boxes(n, 9, 1024.0)).cuda()
    s1 = torch.from_numpy(synth.unique_scores(n, 9)).cuda()
    t = best(lambda: rg.nms_rotated(b1, s1, 0.1))
    res['nms_rotated_single_segment_30000'] = {'boxes': n, 'ms': 1e3 * t, 'boxes_per_s': n / t}
    # C3 (iii): the assigner's reductions of the 8192 x 8192 matrix without the matrix
    a3, b3 = synth.c3_iou(0)
    a3d, b3d = torch.from_numpy(a3).cuda(), torch.from_numpy(b3).cuda()
    t = best(lambda: rg.rbbox_overlaps_max(a3d, b3d))
    res['c3_rowcol_max_8192x8192_fused'] = {'pairs': a3.shape[0] * b3.shape[0], 'ms': 1e3 * t,
                                            'gpairs_per_s': a3.shape[0] * b3.shape[0] / t / 1e9}
    # C3, dense variant (extent 1024: 2.6 % of the pairs overlap, the clipper matters), matrix out
    a4, b4 = synth.c3_iou(0, extent=1024.0)
    a4d, b4d = torch.from_numpy(a4).cuda(), torch.from_numpy(b4).cuda()
    t = best(lambda: rg.rbbox_overlaps(a4d, b4d))
    res['c3_dense_extent1024_8192x8192_python_call'] = {'pairs': a4.shape[0] * b4.shape[0], 'ms': 1e3 * t,
                                                        'gpairs_per_s': a4.shape[0] * b4.shape[0] / t / 1e9}
    del a4d, b4d
    anchors = torch.from_numpy(synth.le90_boxes(196416, 5, 1024.0)).cuda()
    gtb = torch.from_numpy(synth.le90_boxes(100, 6, 1024.0)).cuda()
    asg = rg.MaxIoUAssigner(0.5, 0.4, min_pos_iou=0.0)
    t = best(lambda: asg.assign(anchors, gtb))
    res['retinanet_max_iou_assign_100gt_x_196416_anchors'] = {'pairs': 100 * 196416, 'ms': 1e3 * t,
                                                              'gpairs_per_s': 100 * 196416 / t / 1e9}
    # C5: ~200k detections of 400 overlapping patches, 15 Zipf-skewed classes
    patches, offs, shape = synth.c5_patch_results(0)
    nd = sum(a.shape[0] for p in patches for a in p)
    ts = []
    for _ in range(3):
        cp = [[a.copy() for a in p] for p in patches]     # merge_results shifts its input in place
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        rg.merge_results(cp, offs, shape, 0.1)
        ts.append(time.perf_counter() - t0)
    res['c5_merge_results_200k_host_lists'] = {'boxes': nd, 'ms': 1e3 * min(ts), 'boxes_per_s': nd / min(ts)}
    # `cp` was shifted to image coordinates in place by the last merge_results call; one GROUP per class
    per_cls = [np.concatenate([p[c] for p in cp], axis=0) for c in range(len(cp[0]))]
    counts = [a.shape[0] for a in per_cls]
    dd = torch.from_numpy(np.concatenate(per_cls, axis=0)).cuda()
    from orientedobjectdetection_b200 import ops
    t = best(lambda: ops._nms_grouped(dd[:, :5], dd[:, 5], 0.1, group_sizes=counts), 3)
    res['c5_class_batched_nms_200k_device_resident'] = {'boxes': nd, 'ms': 1e3 * t, 'boxes_per_s': nd / t}
    return res


def secondary_sharded(torch, dist, rg, synth, dev):
    """N > 1: the workloads that SHARD (BASELINE C3 as ONE matrix row-sharded, C4 images over ranks, C5
    classes over ranks).  Every rank holds only its own shard on the device; results reach all ranks
    through peer memory (orientedobjectdetection_b200/dist.py).  Per entry: device_ms = CUDA events
    around the call on the launching stream, ms = wall clock of the call incl. its one host read; both
    the max over ranks, best of a few repetitions."""
    from orientedobjectdetection_b200 import dist as rdist
    import copy
    rank, world = dist.get_rank(), dist.get_world_size()
    res = {'peer_memory': bool(rdist.peer.available())}

    def timed(fn, reps=5):
        fn()
        ev, wl = [], []
        for _ in range(reps):
            torch.cuda.synchronize()
            dist.barrier()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            t0 = time.perf_counter()
            e0.record()
            fn()
            e1.record()
            torch.cuda.synchronize()
            t = torch.tensor([e0.elapsed_time(e1), 1e3 * (time.perf_counter() - t0)], dtype=torch.float64, device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ev.append(float(t[0]))
            wl.append(float(t[1]))
        return min(ev), min(wl)

    def entry(units, key, ev, wl):
        return {key: units, 'device_ms': ev, 'ms': wl, key + '_per_s': units / (wl * 1e-3)}

    # C4: 64 images, each rank uploads only the images it owns
    c4 = [synth.c4_rpn_image(s) for s in range(64)]
    sizes = [int(b.shape[0]) for b, _, _ in c4]
    owner = rdist.lpt_assign([float(v) ** 2 for v in sizes], world)
    mine = [i for i in range(64) if owner[i] == rank]
    bat = rdist.Batched(*[torch.from_numpy(np.concatenate([c4[i][k] for i in mine])).to(dev) for k in range(3)], sizes)
    ev, wl = timed(lambda: rdist.nms_images_sharded(bat, 0.8, 2000))
    res['c4_64img_sharded_over_ranks'] = entry(sum(sizes), 'boxes', ev, wl)
    res['c4_64img_sharded_over_ranks']['input'] = 'each rank holds the proposals of ITS images back to back on the device'
    del bat
    # C5: ~200k detections, classes over ranks; device-resident (each rank holds its classes) ...
    patches, offs, shape = synth.c5_patch_results(0)
    nd = sum(a.shape[0] for p in patches for a in p)
    shifted = copy.deepcopy(patches)
    for p, o in zip(shifted, offs):
        for a in p:
            a[:, :2] += o
    per_cls = [np.concatenate([p[c] for p in shifted], axis=0) for c in range(len(shifted[0]))]
    n_c = [a.shape[0] for a in per_cls]
    own = rdist.lpt_assign([float(v) ** 2 for v in n_c], world)
    dets = [torch.from_numpy(per_cls[c]).to(dev) if own[c] == rank else None for c in range(len(n_c))]
    ev, wl = timed(lambda: rdist.merge_dets_sharded(dets, n_c, 0.1))
    res['c5_200k_class_sharded_device_resident'] = entry(nd, 'boxes', ev, wl)
    del dets
    # ... and through the reference's host-list API (merge_results: numpy lists in, numpy lists out)
    ts = []
    for _ in range(2):
        cp = copy.deepcopy(patches)
        torch.cuda.synchronize()
        dist.barrier()
        t0 = time.perf_counter()
        rdist.merge_results_sharded(cp, offs, shape, 0.1)
        tt = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device=dev)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        ts.append(float(tt))
    res['c5_merge_results_classes_sharded_over_ranks'] = {'boxes': nd, 'ms': 1e3 * min(ts), 'boxes_per_s': nd / min(ts)}
    # C3 as ONE 8192 x 8192 matrix, rows over ranks: (i) compute only, (ii) + the full matrix on every rank,
    # (iii) fused row / column maxima on every rank instead of the matrix
    a3, b3 = synth.c3_iou(0)
    ad, bd = torch.from_numpy(a3).to(dev), torch.from_numpy(b3).to(dev)
    pairs = a3.shape[0] * b3.shape[0]
    ev, wl = timed(lambda: rdist.rbbox_overlaps_sharded(ad, bd, gather=False))
    res['c3_one_8192x8192_matrix_row_sharded_compute_only'] = entry(pairs, 'pairs', ev, wl)
    ev, wl = timed(lambda: rdist.rbbox_overlaps_sharded(ad, bd))
    res['c3_one_8192x8192_matrix_row_sharded_plus_allgather'] = entry(pairs, 'pairs', ev, wl)
    ev, wl = timed(lambda: rdist.rbbox_overlaps_max_sharded(ad, bd))
    res['c3_one_8192x8192_matrix_row_sharded_fused_maxima'] = entry(pairs, 'pairs', ev, wl)
    return res


def fp32_peak():
    """measured FFMA peak of this part (tools/fp32_peak.cu, committed result), else the theoretical one"""
    p = os.path.join(ROOT, 'profiles', 'r02_fp32_ffma_peak.json')
    try:
        d = json.load(open(p))
        return max(r['tflops'] for r in d['runs']), 'measured (profiles/r02_fp32_ffma_peak.json, 16 FFMA chains per thread)'
    except Exception:
        return FP32_PEAK_TFLOPS, 'theoretical at max boost'


def dense_roofline(torch, L, _lib, synth, flush, steps):
    """SURVEY C3's second variant: extent 1024, 2.6 % of the pairs overlap, so the clipper's FP32 work
    is no longer hidden behind the write stream.  The tile kernel alone, CUDA events, L2 flushed."""
    a, b = synth.c3_iou(0, N_ROWS, N_COLS, 1024.0)
    ad, bd = torch.from_numpy(a).cuda(), torch.from_numpy(b).cuda()
    out = torch.empty((N_ROWS, N_COLS), dtype=torch.float32, device='cuda')
    nbytes = L.rotgeo_iou_workspace_bytes(N_ROWS, N_COLS)
    ws = torch.empty(nbytes, dtype=torch.uint8, device='cuda')
    stream = torch.cuda.current_stream().cuda_stream
    _lib.check(L.rotgeo_iou_pairwise(ad.data_ptr(), N_ROWS, 5, bd.data_ptr(), N_COLS, 5, 0, 1e-3, out.data_ptr(),
                                     N_COLS, ws.data_ptr(), nbytes, stream), 'rotgeo_iou_pairwise')
    pk1 = ws.data_ptr()
    pk2 = pk1 + (L.rotgeo_packed_bytes(N_ROWS) + 255) // 256 * 256
    kms = timed_steps(torch, lambda: _lib.check(L.rotgeo_iou_pairwise_packed(pk1, N_ROWS, pk2, N_COLS, 0, out.data_ptr(),
                                                                           N_COLS, stream), 'packed'), steps, flush, None)
    k_ms = float(np.mean(kms))
    nz = int((out > 0).sum().item())
    alg_bytes = 4.0 * N_ROWS * N_COLS + 20.0 * (N_ROWS + N_COLS)
    peak, src = peaks()
    fpk, fsrc = fp32_peak()
    pairs_s = N_ROWS * N_COLS / (k_ms * 1e-3)
    return {'workload': 'C3 dense variant: extent 1024, %d of %d pairs overlap' % (nz, N_ROWS * N_COLS),
            'bound': 'hbm', 'kernel': 'iou_tile_kernel<0,true>', 'kernel_ms': k_ms,
            'achieved': alg_bytes / (k_ms * 1e-3) / 1e9, 'peak': peak, 'unit': 'GB/s',
            'frac': alg_bytes / (k_ms * 1e-3) / 1e9 / peak, 'peak_source': src, 'traffic': None,
            'alternative_bound_fp32': {'flop_per_pair': FLOP_PER_PAIR, 'achieved_tflops': pairs_s * FLOP_PER_PAIR / 1e12,
                                       'peak_tflops': fpk, 'peak_source': fsrc,
                                       'frac': pairs_s * FLOP_PER_PAIR / 1e12 / fpk,
                                       'note': 'the 450 flop/pair model is the reference algorithm\'s mandatory work; '
                                               'the cull skips it for disjoint pairs, so this is not a utilisation'}}


def nms_block(torch, rg, synth):
    """the rotated-NMS half of the metric in the driver-run line: BASELINE C2 (2000 proposals x 15 classes)
    through the public call, device time by CUDA events, per-kernel times by CUPTI, the roofline of the
    suppression-mask pass (B_nms = 24 N + 8 N ceil(N / 64) / 2 + 8 K bytes per
    segment) and the oracle on the same input on one host core (what upstream mmcv's CPU op is)."""
    from torch.profiler import profile, ProfilerActivity
    from oracle import ref_api
    peak, src = peaks()
    out = {}
    for name, dense in (('c2_sparse', False), ('c2_dense', True)):
        mb, ms = synth.c2_multiclass(0, dense=dense)
        mbd, msd = torch.from_numpy(mb).cuda(), torch.from_numpy(ms).cuda()
        per_cls = (ms[:, :-1] > 0.05).sum(0).astype(np.int64)
        ncand = int(per_cls.sum())

        def call():
            return rg.multiclass_nms_rotated(mbd, msd, 0.05, dict(iou_thr=0.1), 2000)
        for _ in range(5):
            kept = call()[0].shape[0]
        torch.cuda.synchronize()
        reps = 20
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            call()
        e1.record()
        e1.synchronize()
        dev_ms = e0.elapsed_time(e1) / reps       # includes the call's own host work between launches
        with profile(activities=[ProfilerActivity.CUDA]) as prof:
            for _ in range(10):
                call()
            torch.cuda.synchronize()
        kt = {}
        for e in prof.key_averages():
            if 'rotgeo::' in e.key:
                short = e.key.split('rotgeo::')[1].split('(')[0].split('<')[0]
                kt[short] = kt.get(short, 0.0) + e.device_time_total / 10.0
        alg = float(sum(24 * n + 8 * n * ((n + 63) // 64) / 2 for n in per_cls) + 8 * kept)
        pairs = float(sum(n * (n - 1) / 2 for n in per_cls))
        mask_us = kt.get('nms_mask_kernel', float('nan'))
        t0 = time.perf_counter()
        ref_api.multiclass_nms_rotated(torch.from_numpy(mb), torch.from_numpy(ms), 0.05, 0.1, 2000)
        cpu_s = time.perf_counter() - t0
        out[name] = {
            'boxes': ncand, 'kept': int(kept), 'ms_per_call': dev_ms, 'boxes_per_s': ncand / (dev_ms * 1e-3),
            'kernels_us': {k: round(v, 2) for k, v in sorted(kt.items())},
            'roofline': {'bound': 'hbm', 'kernel': 'nms_mask_kernel', 'kernel_us': mask_us,
                         'algorithmic_bytes_per_launch': alg, 'achieved': alg / (mask_us * 1e-6) / 1e9, 'peak': peak,
                         'unit': 'GB/s', 'frac': alg / (mask_us * 1e-6) / 1e9 / peak, 'peak_source': src,
                         'pairs_per_launch': pairs, 'gpairs_per_s': pairs / (mask_us * 1e-6) / 1e9, 'traffic': None,
                         'note': 'latency/issue-bound at this size: 15 segments of <= 2000 boxes are 120 tiles'},
            'cpu_baseline': {'value': ncand / cpu_s, 'unit': 'boxes/s', 'cores': 1, 'kind': 'port', 'seconds': cpu_s,
                             'sample': 'the whole call on the same input, 1 repetition'}}
    return out


def measured_traffic():
    """DRAM bytes per launch of the dominant kernel from the newest committed ncu --set full capture
    (profiles/r02_iou_tile_kernel_traffic.json, written from the .ncu-rep by tools/ncu_traffic.py)"""
    for name in ('r02_iou_tile_kernel_traffic.json', 'r01_iou_tile_kernel_traffic.json'):
        p = os.path.join(ROOT, 'profiles', name)
        if os.path.exists(p):
            try:
                return float(json.load(open(p))['dram_bytes_per_launch'])
            except Exception:
                continue
    return None


def run_rotgeo(args, rank, world, local_rank):
    import torch
    import torch.distributed as dist

    import orientedobjectdetection_b200 as rg
    from orientedobjectdetection_b200 import _lib, synth
    assert torch.cuda.is_available(), 'bench.py needs a CUDA device (no CPU fallback)'
    torch.cuda.set_device(local_rank)
    dev = torch.device('cuda', local_rank)
    if world > 1:
        # stdout carries the one JSON line: NCCL's own messages (NCCL_DEBUG=VERSION on some boxes) go to stderr
        os.environ.setdefault('NCCL_DEBUG_FILE', '/dev/stderr')
        dist.init_process_group('nccl', device_id=dev)
    # N > 1: every rank's pinned staging on the NUMA node of its own GPU (8 ranks x 268 MB device->host per step
    # do not fit through one socket's memory: tools/numa_probe.py); before the first pinned allocation
    host_cpus = None
    if world > 1:
        from orientedobjectdetection_b200 import peer as _peer
        host_cpus = _peer.bind_host_to_gpu(dev)
    L = _lib.load()

    # rank r owns rows [r*N_ROWS, (r+1)*N_ROWS) of the global matrix; columns replicated
    a, b = synth.c3_iou(0, N_ROWS, N_COLS, EXTENT)      # the reference arm's block (rank 0 times exactly this one)
    if rank > 0:
        a = synth.le90_boxes(N_ROWS, 1000 + rank, EXTENT)
    a_host, b_host = torch.from_numpy(a).pin_memory(), torch.from_numpy(b).pin_memory()
    a_dev, b_dev = a_host.to(dev), b_host.to(dev)
    out = torch.empty((N_ROWS, N_COLS), dtype=torch.float32, device=dev)
    nbytes = L.rotgeo_iou_workspace_bytes(N_ROWS, N_COLS)
    ws = torch.empty(nbytes, dtype=torch.uint8, device=dev)
    flush_buf = torch.empty(512 << 20, dtype=torch.uint8, device=dev)
    stream = torch.cuda.current_stream().cuda_stream
    pk1 = ws.data_ptr()
    pk2 = pk1 + (L.rotgeo_packed_bytes(N_ROWS) + 255) // 256 * 256

    def flush():
        flush_buf.zero_()

    def step():      # the C-ABI call rbbox_overlaps makes: pack (both sets, clamp 1e-3 fused) + tile kernel
        _lib.check(L.rotgeo_iou_pairwise(a_dev.data_ptr(), N_ROWS, 5, b_dev.data_ptr(), N_COLS, 5, 0,
                                         1e-3, out.data_ptr(), N_COLS, ws.data_ptr(), nbytes, stream),
                   'rotgeo_iou_pairwise')

    def kernel_only():
        _lib.check(L.rotgeo_iou_pairwise_packed(pk1, N_ROWS, pk2, N_COLS, 0, out.data_ptr(), N_COLS,
                                                stream), 'rotgeo_iou_pairwise_packed')

    for _ in range(max(args.warmup, 3)):
        step()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    # The timed region is ~1 ms long and nvidia-smi samples every 100 ms: run the same step
    # untimed for 0.5 s first (all ranks, so that they reach the timed region together), which
    # puts samples taken under this very load on both sides of the timed region.
    t_load = time.perf_counter()
    while time.perf_counter() - t_load < 0.5:
        for _ in range(50):
            step()
        torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    ms = timed_steps(torch, step, args.steps, flush, None)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    total_ms = torch.tensor([sum(ms)], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(total_ms, op=dist.ReduceOp.MAX)
    total_ms = float(total_ms.item())
    ms_per_step = total_ms / args.steps
    value = world * N_ROWS * N_COLS / (ms_per_step * 1e-3) / 1e9

    # dominant kernel alone (same stream, CUDA events, L2 flushed between launches)
    kms = timed_steps(torch, kernel_only, args.steps, flush, None)
    clocks = sampler.stop() if rank == 0 else None
    k_ms = float(np.mean(kms))
    alg_bytes = 4.0 * N_ROWS * N_COLS + 20.0 * (N_ROWS + N_COLS)   # fp32 matrix out + AoS boxes in
    peak, peak_src = peaks()
    achieved = alg_bytes / (k_ms * 1e-3) / 1e9
    pairs_s = N_ROWS * N_COLS / (k_ms * 1e-3)

    # end to end through the public API with HOST buffers (pinned in, pinned out)
    e2e_steps = max(3, min(args.steps, 10))
    # warm-up with two results alive: the pinned-host caching allocator then owns the two
    # 268 MB buffers the steady state alternates between (a cold cudaHostAlloc is ~100 ms)
    keep = [rg.rbbox_overlaps(a_host, b_host), rg.rbbox_overlaps(a_host, b_host)]
    del keep
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        res = rg.rbbox_overlaps(a_host, b_host)
    torch.cuda.synchronize()
    e2e_t = torch.tensor([(time.perf_counter() - t0) / e2e_steps], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(e2e_t, op=dist.ReduceOp.MAX)
    e2e_val = world * N_ROWS * N_COLS / float(e2e_t.item()) / 1e9
    assert res.shape == (N_ROWS, N_COLS) and not res.is_cuda

    line = {
        'metric': 'rotated_iou_gpairs_per_s', 'value': value, 'unit': 'Gpairs/s', 'n_gpus': world,
        'steps': args.steps, 'warmup': max(args.warmup, 3), 'ms_per_step': ms_per_step,
        'higher_is_better': True, 'scaling': 'weak',
        'collective': 'none',    # every rank times its own block: NOT a sharding result (see secondary for those)
        'vs_baseline': None, 'dtype': 'f32',
        'data': 'synthetic',
        'config': {'workload': WORKLOAD,
                   'rows_per_gpu': N_ROWS, 'cols': N_COLS,
                   'l2': 'output 268 MB/step exceeds the 126 MB L2; a 512 MB memset flushes L2 '
                         'between steps, outside the CUDA events',
                   'timing': 'sum of per-step CUDA-event times on the launching stream, max over ranks'},
        'gpu_launches': 2 * args.steps,   # pack_kernel (both sets) + iou_tile_kernel per step
        'e2e': {'value': e2e_val, 'unit': 'Gpairs/s',
                'h2d_bytes_per_step': int(a_host.numel() * 4 + b_host.numel() * 4),
                'd2h_bytes_per_step': int(N_ROWS * N_COLS * 4), 'steps': e2e_steps,
                'api': 'orientedobjectdetection_b200.rbbox_overlaps(pinned CPU tensors) -> pinned CPU tensor '
                       '(four row chunks, copies overlapped with the kernels)',
                'host_binding': ('rank pinned to the %d CPUs NVML lists as local to its GPU' % len(host_cpus)) if host_cpus
                                else 'none'},
        'roofline': {'bound': 'hbm', 'achieved': achieved, 'peak': peak, 'unit': 'GB/s',
                     'frac': achieved / peak, 'traffic': measured_traffic(),
                     'traffic_source': 'committed ncu --set full capture of this kernel (profiles/r02_iou_tile_kernel_traffic.json, round 2), not re-measured in this run',
                     'kernel': 'iou_tile_kernel<0,true>',
                     'kernel_ms': k_ms, 'peak_source': peak_src,
                     'algorithmic_bytes_per_launch': alg_bytes,
                     'fp32_view': {'flop_per_pair': FLOP_PER_PAIR,
                                   'achieved_tflops': pairs_s * FLOP_PER_PAIR / 1e12,
                                   'peak_tflops_theoretical': FP32_PEAK_TFLOPS,
                                   'frac': pairs_s * FLOP_PER_PAIR / 1e12 / FP32_PEAK_TFLOPS}},
        'clocks': clocks,
    }
    if rank == 0 and world == 1 and not args.no_secondary:
        line['roofline_dense'] = dense_roofline(torch, L, _lib, synth, flush, args.steps)
        line['nms'] = nms_block(torch, rg, synth)
        line['secondary'] = secondary_nms(torch, rg, synth)
    if world > 1 and not args.no_secondary:
        try:
            sec = secondary_sharded(torch, dist, rg, synth, dev)
            if rank == 0:
                line['secondary'] = sec
        except Exception as exc:      # the headline line must survive a failure here
            if rank == 0:
                line['secondary'] = {'error': repr(exc)[:300]}
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        ac, bc = a.copy(), b.copy()
        ac[:, 2:4] = np.maximum(ac[:, 2:4], 1e-3)
        bc[:, 2:4] = np.maximum(bc[:, 2:4], 1e-3)
        from oracle import ref
        threads = min(ref.max_threads(), len(os.sched_getaffinity(0)))
        # ~6 s single thread (what upstream mmcv does) + ~10 s on all cores
        v1, t1, r1 = cpu_iou_sample(ac, bc, N_ROWS, 1, 1.0)
        vN, tN, rN = cpu_iou_sample(ac, bc, N_ROWS, threads, 10.0)
        line['cpu_baseline'] = {
            'value': vN, 'unit': 'Gpairs/s', 'cores': threads, 'kind': 'port',
            'sample': 'the whole 8192x8192 C3 block, %d repetitions, %.1f s, best' % (rN, tN),
            'single_thread': {'value': v1, 'cores': 1,
                              'sample': 'the whole 8192x8192 C3 block, %d repetition(s), %.1f s' % (r1, t1)}}
    if rank == 0:
        emit(line)
    if world > 1:
        dist.destroy_process_group()


_JSON_OUT = None


def emit(obj):
    """the one JSON line, on the process's ORIGINAL stdout"""
    out = _JSON_OUT or sys.stdout
    out.write(json.dumps(obj) + '\n')
    out.flush()


def main():
    global _JSON_OUT
    args = parse()
    # stdout carries exactly one JSON line.  Libraries write there too (NCCL prints its version
    # banner to stdout when the box exports NCCL_DEBUG=VERSION, and NCCL_DEBUG_FILE does not move
    # it): keep the original stdout for the JSON line and point fd 1 at stderr for everything else.
    sys.stdout.flush()
    _JSON_OUT = os.fdopen(os.dup(1), 'w')
    os.dup2(2, 1)
    rank = int(os.environ.get('RANK', 0))
    world = int(os.environ.get('WORLD_SIZE', 1))
    local_rank = int(os.environ.get('LOCAL_RANK', 0))
    if args.impl == 'reference':
        run_reference(args, rank)
    else:
        run_rotgeo(args, rank, world, local_rank)


if __name__ == '__main__':
    main()
