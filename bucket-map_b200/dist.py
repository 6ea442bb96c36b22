"""Multi-GPU plumbing: one process per GPU, reads sharded contiguously, index replicated.

torch.distributed is used exactly twice on the path: one broadcast of the q-gram bit matrix after it has been built
or loaded on rank 0, and one variable-length gather of the fixed-size location records to rank 0.  No collective sits
inside the per-read work, so there is nothing to fuse.  Works with the nccl backend (device tensors that alias the
library's buffers) and with gloo (CPU tensors; used by the world_size-2 CPU tests).

Since round 2 the library itself replicates the index and gathers the records (csrc/bm_comm.cu: bm_bcast_index, bm_bcast_ref,
bm_index_build_sharded, bm_gather_results) and bench.py uses those; this module stays as the torch.distributed form of the
same plumbing for callers that already run a process group, and as what the gloo tests exercise on machines without a GPU.
"""
import contextlib
import os
import threading
import time

import numpy as np
import torch
import torch.distributed as dist

from . import LOC_DTYPE


TRACE = {} if os.environ.get("BM_GATHER_TRACE") else None     # per-phase host milliseconds of LocationGatherer.gather


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
    rank order is read order; shards may differ in size).  Every rank copies its records into a persistent send buffer
    and sends dst a two-word header (records, reads); dst reads the headers -- its only host synchronisation, the
    other ranks have none -- and receives each rank's per-read counts and records directly at their final offsets in
    the result, so there is no padding and no compaction pass.  Point-to-point sends / receives: NCCL over NVLink when
    the tensors are on the GPU, gloo on CPU tensors in the tests.  A prefix sum gives loc_off; on CUDA the result is
    copied into pinned host memory on a side stream (wait() before reading it; host_copy=False leaves it on dst's
    device).  `dst` is a global rank.

    gather() runs the exchange on the calling thread.  gather_async() stages the records (after which the library may
    overwrite its result buffers) and hands the exchange to a worker thread; wait() joins it.  Note that on the device
    the exchange does not overlap a running mapper kernel, whatever the stream priorities: the mapper fills every SM
    and the small NCCL / copy kernels are only dispatched once its grid is exhausted (measured).
    Collectives are issued by one thread at a time, in call order, on every rank."""

    def __init__(self, dst=0, group=None, host_copy=True):
        self.dst, self.group, self.host_copy = dst, group, host_copy
        self.world, self.rank = dist.get_world_size(group), dist.get_rank(group)
        self.send = self.cnt_send = None
        self.side = None
        self.h_off = self.h_locs = None
        self.result = None
        self.worker = None
        self.worker_error = None
        self.hp = None                # high-priority CUDA stream of the staging copies and the exchange

    def _stream(self, device):
        """Everything the gatherer enqueues runs on one high-priority stream (and NCCL's stream is created with high
        priority, see init_process_group in bench.py): the block scheduler otherwise dispatches the small copy / NCCL /
        compaction kernels only after the whole grid of a mapper kernel that was launched before them."""
        if device.type != "cuda":
            return contextlib.nullcontext()
        if self.hp is None:
            self.hp = torch.cuda.Stream(device, priority=-1)
        return torch.cuda.stream(self.hp)

    # -- step 1, on the caller's thread: the rank's records into persistent send buffers ------------------------------
    def _stage(self, counts, locs_bytes):
        device, rec, n = counts.device, locs_bytes.shape[1], locs_bytes.shape[0]
        if self.send is None or self.send.device != device or self.send.shape[0] < n or self.send.shape[1] != rec:
            self.send = torch.zeros((int(n * 1.25) + 1024, rec), dtype=torch.uint8, device=device)
        if self.cnt_send is None or self.cnt_send.device != device or self.cnt_send.shape[0] != counts.shape[0]:
            self.cnt_send = torch.empty(counts.shape[0], dtype=torch.int32, device=device)
        with self._stream(device):
            if device.type == "cuda":
                self.hp.wait_stream(torch.cuda.current_stream(device))     # counts / locs_bytes were produced there
            self.send[:n] = locs_bytes
            self.cnt_send.copy_(counts)
            if device.type == "cuda":
                counts.record_stream(self.hp); locs_bytes.record_stream(self.hp)
                self.hp.synchronize()                  # the caller's buffers may be reused from here on
        return n

    # -- step 2: headers to dst, then every rank's counts and records straight into their place in dst's result ----------
    def _exchange(self, n):
        with self._stream(self.send.device):
            return self._exchange_on_stream(n)

    def _p2p(self, ops):
        if ops:
            for w in dist.batch_isend_irecv(ops):
                w.wait()                         # stream-level on NCCL, blocking on gloo

    def _exchange_on_stream(self, n):
        device, rec = self.send.device, self.send.shape[1]
        n_reads = self.cnt_send.shape[0]
        t0 = time.perf_counter()
        hdr = torch.tensor([n, n_reads], dtype=torch.int64, device=device)
        if self.rank != self.dst:
            # no host synchronisation on this side: three sends are enqueued and the rank goes on with its next batch
            self._p2p([dist.P2POp(dist.isend, hdr, self.dst, self.group)])
            ops = [dist.P2POp(dist.isend, self.cnt_send, self.dst, self.group)] if n_reads else []
            if n:
                ops.append(dist.P2POp(dist.isend, self.send[:n], self.dst, self.group))
            self._p2p(ops)
            self.result = None
            return None
        hdrs = torch.zeros((self.world, 2), dtype=torch.int64, device=device)
        hdrs[self.dst] = hdr
        self._p2p([dist.P2POp(dist.irecv, hdrs[r], r, self.group) for r in range(self.world) if r != self.dst])
        h = hdrs.tolist()                        # the one host synchronisation of the exchange (dst only)
        t1 = time.perf_counter()
        sizes, reads = [int(x[0]) for x in h], [int(x[1]) for x in h]
        total, total_reads = sum(sizes), sum(reads)
        cnt = torch.empty(total_reads, dtype=torch.int32, device=device)
        locs = torch.empty((total, rec), dtype=torch.uint8, device=device)
        ops, o, l = [], 0, 0
        for r in range(self.world):
            if r == self.dst:
                cnt[o:o + reads[r]] = self.cnt_send
                locs[l:l + sizes[r]] = self.send[:n]
            else:
                if reads[r]:
                    ops.append(dist.P2POp(dist.irecv, cnt[o:o + reads[r]], r, self.group))
                if sizes[r]:
                    ops.append(dist.P2POp(dist.irecv, locs[l:l + sizes[r]], r, self.group))
            o += reads[r]; l += sizes[r]
        self._p2p(ops)
        off = torch.zeros(total_reads + 1, dtype=torch.int64, device=device)
        torch.cumsum(cnt, 0, out=off[1:])
        if TRACE is not None:
            t2 = time.perf_counter()
            for k, v in (("headers", t1 - t0), ("enqueue", t2 - t1)):
                TRACE[k] = TRACE.get(k, 0.0) + v * 1e3
            TRACE["n"] = TRACE.get("n", 0) + 1
        if device.type == "cuda" and not self.host_copy:
            self.result = (off, locs, False)       # results stay on dst's device, like a single GPU's resident results
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
            self.result = (self.h_off[:off.numel()], self.h_locs[:total], True)
        else:
            self.result = (off, locs, True)
        return self.result[:2]

    def gather(self, counts, locs_bytes):
        """counts: int32 [n_reads_local] (equal n_reads_local on every rank, or use gather_variable-based
        gather_locations); locs_bytes: uint8 [n_locs_local, rec].  Returns (loc_off, locs) handles on dst -- device
        tensors, or their pinned host copies filled asynchronously -- and None elsewhere."""
        self._join()
        return self._exchange(self._stage(counts, locs_bytes))

    def gather_async(self, counts, locs_bytes):
        """as gather(), with the exchange on a worker thread; returns once the records are staged.  wait() (or the next
        gather) joins the worker."""
        self._join()
        n = self._stage(counts, locs_bytes)
        device = self.send.device

        def run():
            try:
                if device.type == "cuda":
                    torch.cuda.set_device(device)
                self._exchange(n)
            except BaseException as e:      # surfaced by _join on the caller's thread
                self.worker_error = e

        self.worker = threading.Thread(target=run, name="bm-gather", daemon=True)
        self.worker.start()

    def _join(self):
        if self.worker is not None:
            self.worker.join()
            self.worker = None
        if self.worker_error is not None:
            e, self.worker_error = self.worker_error, None
            raise e

    def wait(self):
        """joins a pending gather_async; on dst returns the host arrays (loc_off uint32[n_total+1], locs LOC_DTYPE[n]) of
        the last gather when it kept a host copy, else None"""
        self._join()
        if self.hp is not None:
            self.hp.synchronize()                     # device-resident results are complete from here on
        if self.result is None or not self.result[2]:
            return None
        if self.hp is not None:
            self.hp.synchronize()
        if self.side is not None:
            self.side.synchronize()
        off, locs = self.result[:2]
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
