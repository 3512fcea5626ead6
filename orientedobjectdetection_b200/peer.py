"""Peer-memory windows for the multi-GPU exchange (one process per GPU, NVLink / NVSwitch).

The data path is librotgeo's own kernels (csrc/rotgeo_dist.cu: barrier by release/acquire
flags in the peers' memory, sparse push of a row-sharded IoU matrix, key exchange for the
fused maxima, keep-list push).  PyTorch is plumbing here: ``torch.distributed._symmetric_memory``
allocates one buffer per rank and maps every rank's buffer into every process; the library
only ever sees the table of peer base addresses.

The reference has no multi-GPU form of this path (python loops inside one rank, reference
mmrotate/models/dense_heads/rotated_rpn_head.py:414-426,515, mmrotate/core/patch/merge_results.py:90-119).
"""
import ctypes
import importlib

import torch
import torch.distributed as dist

from . import _lib

_WINDOWS = {}


def available(group=None):
    """peer windows need an initialised NCCL group on CUDA devices and torch symmetric memory"""
    if not (dist.is_available() and dist.is_initialized()):
        return False
    if dist.get_backend(group) != 'nccl' or not torch.cuda.is_available():
        return False
    try:
        importlib.import_module('torch.distributed._symmetric_memory')
    except Exception:
        return False
    return dist.get_world_size(group) <= 16


class PeerWindow(object):
    """``nbytes`` of symmetric device memory per rank (the first 4 KB are the signal pad of the
    library's barrier) + the table of peer addresses + the epoch counter of the exchange calls.
    Collective: every rank of ``group`` must construct it at the same point."""

    def __init__(self, nbytes, group=None):
        import torch.distributed._symmetric_memory as symm
        self.group = group if group is not None else dist.group.WORLD
        self.world = dist.get_world_size(self.group)
        self.rank = dist.get_rank(self.group)
        self.dev = torch.device('cuda', torch.cuda.current_device())
        self.lib = _lib.load()
        self.pad = int(self.lib.rotgeo_peer_pad_bytes())
        self.nbytes = int(nbytes)
        self.buf = symm.empty(self.nbytes, dtype=torch.uint8, device=self.dev)
        self.hdl = symm.rendezvous(self.buf, self.group.group_name)
        self.buf[:self.pad].zero_()
        torch.cuda.synchronize(self.dev)
        self.hdl.barrier()                     # every pad is zero before anybody signals
        self.ptrs = (ctypes.c_void_p * self.world)(*[int(p) for p in self.hdl.buffer_ptrs])
        self.epoch = 0

    def take(self, k):
        """first of ``k`` fresh epoch values (all ranks call the same sequence)"""
        e = self.epoch + 1
        self.epoch += k
        return e

    def view(self, offset, shape, dtype):
        """tensor over MY window; valid until the region is written by a later exchange"""
        n = 1
        for s in shape:
            n *= int(s)
        nb = n * torch.empty(0, dtype=dtype).element_size()
        return self.buf[offset:offset + nb].view(dtype).view(*shape)

    def barrier(self, stream):
        _lib.check(self.lib.rotgeo_peer_barrier(self.ptrs, self.world, self.rank, self.take(1), stream),
                   'rotgeo_peer_barrier')

    def check_timeout(self):
        """a barrier that gave up (a peer died) leaves its epoch in slot 64 of the pad"""
        bad = int(self.buf[512:520].view(torch.int64).item())
        if bad:
            raise RuntimeError('rotgeo: peer barrier %d timed out (a rank of the group did not arrive)' % bad)


def window(nbytes, group=None):
    """the cached window of this (group, device), grown when a call needs more.  Growing is a
    collective: all ranks ask for the same sizes in the same order (the sharded entry points
    derive them from arguments every rank shares)."""
    key = (id(group) if group is not None else 0, torch.cuda.current_device())
    w = _WINDOWS.get(key)
    if w is None or w.nbytes < nbytes:
        w = PeerWindow(max(int(nbytes * 1.25), 1 << 20), group)
        _WINDOWS[key] = w
    return w


def release():
    _WINDOWS.clear()


def bind_host_to_gpu(device=None):
    """Pin the calling thread to the CPUs NVML reports as local to ``device`` (default: the current
    CUDA device), so that pinned staging memory allocated afterwards (the result buffers of the
    host-tensor IoU path, the merge staging rows) is first-touched on the GPU's own NUMA node.
    One process per GPU on a two-socket box otherwise puts every rank's pinned buffers — and 8 x 54
    GB/s of device->host copies — on whatever node the launcher started on.  Process-wide side
    effect, so it is a call the application makes (bench.py does at N > 1), never done implicitly.
    Returns the sorted CPU list, or None when NVML or the UUID lookup is not available."""
    import os
    try:
        import pynvml
        idx = torch.cuda.current_device() if device is None else torch.device(device).index
        uuid = str(torch.cuda.get_device_properties(idx).uuid)
        pynvml.nvmlInit()
        handle = pynvml.nvmlDeviceGetHandleByUUID(('GPU-' + uuid) if not uuid.startswith('GPU-') else uuid)
        pynvml.nvmlDeviceSetCpuAffinity(handle)
        return sorted(os.sched_getaffinity(0))
    except Exception:
        return None
