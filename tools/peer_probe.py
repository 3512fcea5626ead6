"""What peer-memory plumbing works on this box (run under torchrun, 2+ ranks)?  (a) torch symmetric
memory, (b) CUDA IPC of a torch storage, (c) NCCL all-gather bandwidth of the C3 matrix for reference.
    python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 tools/peer_probe.py"""
import json, os, sys, time, traceback
import torch
import torch.distributed as dist

rank, world, local = int(os.environ['RANK']), int(os.environ['WORLD_SIZE']), int(os.environ['LOCAL_RANK'])
torch.cuda.set_device(local)
dev = torch.device('cuda', local)
dist.init_process_group('nccl', device_id=dev)
out = {'world': world, 'p2p_0_1': torch.cuda.can_device_access_peer(0, 1) if torch.cuda.device_count() > 1 else None}


def ev_time(fn, reps=10):
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


# (c) NCCL reference
n = 8192 * 8192
full = torch.empty(n, dtype=torch.float32, device=dev)
part = torch.ones(n // world, dtype=torch.float32, device=dev)
ms = ev_time(lambda: dist.all_gather_into_tensor(full, part))
out['nccl_allgather_268MB_ms'] = ms
out['nccl_allgather_recv_GBps_per_gpu'] = (n * 4 * (world - 1) / world) / (ms * 1e-3) / 1e9
small = torch.ones(16384, dtype=torch.int64, device=dev)
sfull = torch.empty(16384 * world, dtype=torch.int64, device=dev)
out['nccl_allgather_128KB_ms'] = ev_time(lambda: dist.all_gather_into_tensor(sfull, small), 50)
out['nccl_barrier_like_allreduce_8B_ms'] = ev_time(lambda: dist.all_reduce(small[:1]), 50)

# (a) symmetric memory
try:
    import torch.distributed._symmetric_memory as symm
    t = symm.empty(1 << 20, dtype=torch.float32, device=dev)
    hdl = symm.rendezvous(t, dist.group.WORLD.group_name)
    out['symm_ok'] = True
    out['symm_ptrs'] = [hex(p) for p in hdl.buffer_ptrs]
    out['symm_signal_pad'] = hdl.signal_pad_size if hasattr(hdl, 'signal_pad_size') else None
    t.fill_(float(rank + 1))
    hdl.barrier()
    peer = hdl.get_buffer((rank + 1) % world, (1 << 20,), torch.float32)
    out['symm_peer_read'] = float(peer[123].item())
    peer[7:9] = 100.0 + rank            # remote write
    torch.cuda.synchronize()
    hdl.barrier()
    out['symm_after_remote_write'] = float(t[7].item())
    out['symm_barrier_ms'] = ev_time(lambda: hdl.barrier(), 50)
    big = symm.empty(n, dtype=torch.float32, device=dev)
    bh = symm.rendezvous(big, dist.group.WORLD.group_name)
    src = torch.ones(n // world, dtype=torch.float32, device=dev)

    def push():
        for r in range(world):
            bh.get_buffer(r, (n,), torch.float32)[rank * (n // world):(rank + 1) * (n // world)].copy_(src)
    out['symm_p2p_push_allgather_268MB_ms'] = ev_time(push)
    out['symm_multicast_ptr'] = hex(bh.multicast_ptr) if getattr(bh, 'multicast_ptr', 0) else None
except Exception as exc:
    out['symm_ok'] = False
    out['symm_error'] = traceback.format_exc()[-600:]

# (b) CUDA IPC of a torch storage
try:
    buf = torch.zeros(1 << 20, dtype=torch.float32, device=dev)
    handle = buf.untyped_storage()._share_cuda_()
    handles = [None] * world
    dist.all_gather_object(handles, handle)
    h = handles[(rank + 1) % world]
    st = torch.UntypedStorage._new_shared_cuda(*h)
    peer = torch.empty(0, dtype=torch.float32, device=st.device).set_(st)
    out['ipc_peer_device'] = str(peer.device)
    # map it on MY device pointer-wise: a tensor of the peer device can be written from a kernel on my device
    peer[3] = 7.0 + rank
    torch.cuda.synchronize(); dist.barrier()
    out['ipc_after_remote_write'] = float(buf[3].item())
    out['ipc_ok'] = True
    out['ipc_ptr'] = hex(peer.data_ptr())
except Exception as exc:
    out['ipc_ok'] = False
    out['ipc_error'] = traceback.format_exc()[-600:]

if rank == 0:
    print(json.dumps(out, indent=1))
dist.destroy_process_group()
