"""Probe (run under torchrun, one rank per GPU): aggregate device->host bandwidth of all ranks copying a
268 MB matrix at the same time into pinned memory, (a) allocated where the launcher put the process,
(b) allocated after peer.bind_host_to_gpu() (CPUs local to the GPU, NVML) — the ceiling of bench.py's
`e2e` at N > 1.
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29511 tools/numa_probe.py"""
import json, os, subprocess, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import torch.distributed as dist
from orientedobjectdetection_b200 import peer

rank, world, local = int(os.environ.get('RANK', 0)), int(os.environ.get('WORLD_SIZE', 1)), int(os.environ.get('LOCAL_RANK', 0))
torch.cuda.set_device(local)
if world > 1:
    dist.init_process_group('nccl', device_id=torch.device('cuda', local))
NBYTES = 8192 * 8192 * 4
src = torch.empty(NBYTES, dtype=torch.uint8, device='cuda')
src.zero_()


def node_of_cpus(cpus):
    nodes = set()
    for n in os.listdir('/sys/devices/system/node') if os.path.isdir('/sys/devices/system/node') else []:
        if n.startswith('node') and os.path.exists('/sys/devices/system/node/%s/cpulist' % n):
            txt = open('/sys/devices/system/node/%s/cpulist' % n).read().strip()
            have = set()
            for part in txt.split(','):
                if '-' in part:
                    a, b = part.split('-'); have |= set(range(int(a), int(b) + 1))
                elif part:
                    have.add(int(part))
            if have & set(cpus):
                nodes.add(n)
    return sorted(nodes)


def measure(extra):
    host = torch.empty(NBYTES + extra, dtype=torch.uint8, pin_memory=True)   # a size the host cache has not seen
    host[:NBYTES].copy_(src, non_blocking=True); torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(10):
        host[:NBYTES].copy_(src, non_blocking=True)
    torch.cuda.synchronize()
    dt = torch.tensor([(time.perf_counter() - t0) / 10], dtype=torch.float64, device='cuda')
    mine = NBYTES / float(dt.item()) / 1e9
    if world > 1:
        dist.all_reduce(dt, op=dist.ReduceOp.MAX)
    return mine, world * NBYTES / float(dt.item()) / 1e9


before_cpus = sorted(os.sched_getaffinity(0))
a_mine, a_all = measure(4096)
cpus = peer.bind_host_to_gpu()
b_mine, b_all = measure(8192)
rec = {'rank': rank, 'cpus_before': len(before_cpus), 'nodes_before': node_of_cpus(before_cpus),
       'cpus_after': len(cpus) if cpus else None, 'nodes_after': node_of_cpus(cpus) if cpus else None,
       'd2h_gbs_unbound': round(a_mine, 1), 'd2h_gbs_bound': round(b_mine, 1)}
recs = [None] * world
if world > 1:
    dist.all_gather_object(recs, rec)
else:
    recs = [rec]
if rank == 0:
    topo = subprocess.run(['nvidia-smi', 'topo', '-m'], capture_output=True, text=True).stdout
    print(json.dumps({'world': world, 'aggregate_d2h_gbs_unbound': round(a_all, 1), 'aggregate_d2h_gbs_bound': round(b_all, 1),
                      'ranks': recs}))
    print(topo)
if world > 1:
    dist.destroy_process_group()
