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


class LocationGatherer:
    """Variable-length gather of per-read location lists to rank `dst`, in global read order (shards are contiguous, so
    rank order is read order).  Buffers persist across calls: one fixed-capacity padded send buffer per rank, one
    [world, cap] receive buffer on dst (grown on demand), so a step costs one tiny all_gather of the record counts, one
    gather of the per-read counts, one gather of the records -- NCCL over NVLink when the tensors are on the GPU, gloo
    on CPU tensors in the tests -- then a device-side prefix sum / compaction and, on CUDA, an asynchronous copy of the
    result into pinned host memory on a side stream (wait() before reading it; host_copy=False leaves the gathered
    records on dst's device)."""

    def __init__(self, dst=0, group=None, host_copy=True):
        self.dst, self.group, self.host_copy = dst, group, host_copy
        self.world, self.rank = dist.get_world_size(group), dist.get_rank(group)
        self.cap = 0
        self.send = self.recv = self.cnt_recv = None
        self.side = None
        self.h_off = self.h_locs = None
        self.result = None

    def _ensure(self, cap, n_reads, device, rec):
        if cap > self.cap or self.send is None or self.send.device != device:
            self.cap = int(cap * 1.25) + 1024
            self.send = torch.zeros((self.cap, rec), dtype=torch.uint8, device=device)
            self.recv = (torch.empty((self.world, self.cap, rec), dtype=torch.uint8, device=device)
                         if self.rank == self.dst else None)
        if self.rank == self.dst and (self.cnt_recv is None or self.cnt_recv.shape[1] != n_reads
                                      or self.cnt_recv.device != device):
            self.cnt_recv = torch.empty((self.world, n_reads), dtype=torch.int32, device=device)

    def gather(self, counts, locs_bytes):
        """counts: int32 [n_reads_local] (equal n_reads_local on every rank, or use gather_variable-based
        gather_locations); locs_bytes: uint8 [n_locs_local, rec].  Returns (loc_off, locs) handles on dst -- device
        tensors whose pinned host copies are filled asynchronously -- and None elsewhere."""
        device, rec = counts.device, locs_bytes.shape[1]
        n = torch.tensor([locs_bytes.shape[0]], dtype=torch.int64, device=device)
        sizes = [torch.zeros_like(n) for _ in range(self.world)]
        dist.all_gather(sizes, n, group=self.group)
        sizes = torch.cat(sizes).tolist()
        self._ensure(max(sizes), counts.shape[0], device, rec)
        self.send[:locs_bytes.shape[0]] = locs_bytes
        if self.rank == self.dst:
            dist.gather(counts, list(self.cnt_recv.unbind(0)), dst=self.dst, group=self.group)
            dist.gather(self.send, list(self.recv.unbind(0)), dst=self.dst, group=self.group)
        else:
            dist.gather(counts, None, dst=self.dst, group=self.group)
            dist.gather(self.send, None, dst=self.dst, group=self.group)
            return None
        total = sum(sizes)
        off = torch.zeros(self.cnt_recv.numel() + 1, dtype=torch.int64, device=device)
        torch.cumsum(self.cnt_recv.view(-1), 0, out=off[1:])
        locs = torch.cat([self.recv[r, :sizes[r]] for r in range(self.world)]) if total else self.recv[0, :0]
        if device.type == "cuda" and not self.host_copy:
            self.result = None                     # results stay on dst's device, like a single GPU's resident results
            return (off, locs)
        if device.type == "cuda":
            if self.side is None:
                self.side = torch.cuda.Stream(device)
            if self.h_off is None or self.h_off.numel() < off.numel():
                self.h_off = torch.empty(off.numel(), dtype=torch.int64, pin_memory=True)
            if self.h_locs is None or self.h_locs.shape[0] < max(total, 1):
                self.h_locs = torch.empty((int(total * 1.25) + 1024, rec), dtype=torch.uint8, pin_memory=True)
            self.side.wait_stream(torch.cuda.current_stream(device))
            with torch.cuda.stream(self.side):
                self.h_off[:off.numel()].copy_(off, non_blocking=True)
                self.h_locs[:total].copy_(locs, non_blocking=True)
                off.record_stream(self.side); locs.record_stream(self.side)
            self.result = (self.h_off[:off.numel()], self.h_locs[:total])
        else:
            self.result = (off, locs)
        return self.result

    def wait(self):
        """host arrays (loc_off uint32[n_total+1], locs LOC_DTYPE[n]) of the last gather on dst"""
        if self.result is None:
            return None
        if self.side is not None:
            self.side.synchronize()
        off, locs = self.result
        return (off.numpy().astype(np.uint32),
                locs.numpy().reshape(-1, LOC_DTYPE.itemsize).view(LOC_DTYPE).reshape(-1))


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
    """one-shot form for shards of unequal size.  counts: int32 tensor [n_reads_local] (locations per read);
    locs_bytes: uint8 tensor [n_locs_local, 20].  On dst returns (loc_off uint32[n_total+1], locs LOC_DTYPE[n]) in
    global read order; None elsewhere."""
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
