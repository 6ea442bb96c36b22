"""Multi-GPU plumbing: one process per GPU, reads sharded contiguously, index replicated.

torch.distributed is used exactly twice on the path: one broadcast of the q-gram bit matrix after it has been built
or loaded on rank 0, and one variable-length gather of the fixed-size location records to rank 0.  No collective sits
inside the per-read work, so there is nothing to fuse.  Works with the nccl backend (device tensors that alias the
library's buffers) and with gloo (CPU tensors; used by the world_size-2 CPU tests).
"""
import numpy as np
import torch
import torch.distributed as dist

from . import LOC_DTYPE


class _DevMem:
    """exposes a raw device pointer through __cuda_array_interface__ so torch can alias it without a copy"""

    def __init__(self, ptr, nbytes):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 2}


def device_bytes(ptr, nbytes, device):
    """uint8 torch tensor over [ptr, ptr+nbytes) on `device` (no copy; the library keeps ownership)"""
    return torch.as_tensor(_DevMem(ptr, nbytes), device=torch.device("cuda", device))


def shard_bounds(n, rank, world):
    """contiguous read shard [lo, hi) of rank `rank` (so rank order == read order when results are concatenated)"""
    return n * rank // world, n * (rank + 1) // world


def broadcast_index(ctx, k2i, q, n_rows, NB, device, src=0, group=None, chunk=1 << 30):
    """replicate the resident bit matrix of rank `src` into every rank's context (ncclBroadcast over NVLink)"""
    if dist.get_rank(group) != src:
        ctx.index_alloc(k2i, q, n_rows, NB)
    ptr, nbytes = ctx.index_device()
    ctx.sync()
    for o in range(0, nbytes, chunk):
        t = device_bytes(ptr + o, min(chunk, nbytes - o), device)
        dist.broadcast(t, src, group=group)
    torch.cuda.synchronize(device)
    return nbytes


def gather_variable(t, dst=0, group=None):
    """gather tensors whose first dimension differs per rank; returns the list on dst, None elsewhere"""
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    n = torch.tensor([t.shape[0]], dtype=torch.int64, device=t.device)
    sizes = [torch.zeros_like(n) for _ in range(world)]
    dist.all_gather(sizes, n, group=group)
    sizes = [int(s.item()) for s in sizes]
    mx = max(max(sizes), 1)
    pad = torch.zeros((mx,) + tuple(t.shape[1:]), dtype=t.dtype, device=t.device)
    pad[:t.shape[0]] = t
    bufs = [torch.empty_like(pad) for _ in range(world)] if rank == dst else None
    dist.gather(pad, bufs, dst=dst, group=group)
    if rank != dst:
        return None
    return [b[:s] for b, s in zip(bufs, sizes)]


def gather_locations(counts, locs_bytes, dst=0, group=None):
    """counts: int32 tensor [n_reads_local] (locations per read); locs_bytes: uint8 tensor [n_locs_local, 20].
    On dst returns (loc_off uint32[n_total+1], locs LOC_DTYPE[n]) in global read order; None elsewhere."""
    c = gather_variable(counts, dst, group)
    l = gather_variable(locs_bytes, dst, group)
    if c is None:
        return None
    cnt = torch.cat(c).cpu().numpy().astype(np.uint64)
    off = np.zeros(len(cnt) + 1, np.uint32)
    off[1:] = np.cumsum(cnt).astype(np.uint32)
    locs = torch.cat(l).cpu().numpy().reshape(-1, LOC_DTYPE.itemsize).view(LOC_DTYPE).reshape(-1)
    return off, locs


def device_results(ctx, device):
    """(counts int32 [n_reads], locs uint8 [n_locs, 20]) aliasing the context's device-resident results"""
    p_off, p_locs, n = ctx.locate_device_buffers()
    off = device_bytes(p_off, (ctx.n_reads + 1) * 4, device).view(torch.int32)
    counts = off[1:] - off[:-1]
    if n:
        locs = device_bytes(p_locs, n * LOC_DTYPE.itemsize, device).view(-1, LOC_DTYPE.itemsize)
    else:
        locs = torch.zeros((0, LOC_DTYPE.itemsize), dtype=torch.uint8, device=torch.device("cuda", device))
    return counts, locs
