"""CPU, world_size 2 over gloo: the host-side multi-GPU plumbing (read sharding, variable-length gather of location
records to rank 0 in global read order).  The index broadcast aliases device memory and is covered on the GPU box."""
import os
import socket

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import bucketmap_b200 as bm
from bucketmap_b200 import dist as bmd


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _fake_results(n_reads_total, rank, world, seed=5):
    """deterministic per-read location lists; rank r owns the contiguous shard shard_bounds(n, r, world)"""
    rng = np.random.default_rng(seed)
    counts = rng.integers(0, 4, n_reads_total)
    locs = np.zeros(int(counts.sum()), bm.LOC_DTYPE)
    locs["bucket"] = rng.integers(0, 1000, len(locs))
    locs["offset"] = rng.integers(-5, 65000, len(locs))
    locs["votes"] = rng.integers(6, 12, len(locs))
    locs["is_original"] = rng.integers(0, 2, len(locs))
    off = np.concatenate([[0], np.cumsum(counts)])
    lo, hi = bmd.shard_bounds(n_reads_total, rank, world)
    return counts, locs, off, lo, hi


def _worker(rank, world, port, n_reads, out_dir):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    counts, locs, off, lo, hi = _fake_results(n_reads, rank, world)
    my_counts = torch.from_numpy(counts[lo:hi].astype(np.int32))
    my_locs = locs[off[lo]:off[hi]]
    my_bytes = torch.from_numpy(my_locs.view(np.uint8).reshape(-1, bm.LOC_DTYPE.itemsize).copy())
    res = bmd.gather_locations(my_counts, my_bytes, dst=0)
    if True:
        # the persistent-buffer gatherer the bench uses gives the same, twice in a row (shards of any size)
        g = bmd.LocationGatherer(dst=0)
        for _ in range(2):
            r2 = g.gather(my_counts, my_bytes)
            assert (r2 is None) == (rank != 0)
        if rank == 0:
            o2, l2 = g.wait()
            assert np.array_equal(o2, res[0]) and np.array_equal(l2, res[1])
        # the overlapped form: records are staged before gather_async returns (the caller's buffers are then scribbled
        # over, as the library does with the next batch); a growing batch re-sizes the agreed capacity
        for rep in (1, 3, 2):
            c_rep, b_rep = my_counts.repeat(rep), my_bytes.repeat(rep, 1)
            g.gather_async(c_rep, b_rep)
            c_rep.zero_(); b_rep.fill_(255)
            got = g.wait()
            assert (got is None) == (rank != 0)
            if rank == 0:
                # expected: every rank's shard repeated rep times, in rank order
                exp_cnt, exp_locs = [], []
                for r in range(world):
                    rlo, rhi = bmd.shard_bounds(n_reads, r, world)
                    exp_cnt.append(np.tile(counts[rlo:rhi], rep))
                    exp_locs.append(np.tile(locs[off[rlo]:off[rhi]], rep))
                exp_off = np.concatenate([[0], np.cumsum(np.concatenate(exp_cnt))]).astype(np.uint32)
                assert np.array_equal(got[0], exp_off) and np.array_equal(got[1], np.concatenate(exp_locs))
    if rank == 0:
        g_off, g_locs = res
        np.save(os.path.join(out_dir, "off.npy"), g_off)
        np.save(os.path.join(out_dir, "locs.npy"), g_locs)
    else:
        assert res is None
    dist.barrier()
    dist.destroy_process_group()


def test_shard_bounds_cover_all_reads():
    for n in (0, 1, 7, 1000, 1_000_003):
        for world in (1, 2, 4, 8):
            b = [bmd.shard_bounds(n, r, world) for r in range(world)]
            assert b[0][0] == 0 and b[-1][1] == n
            assert all(b[i][1] == b[i + 1][0] for i in range(world - 1))
            sizes = [hi - lo for lo, hi in b]
            assert max(sizes) - min(sizes) <= 1


def test_gather_locations_world2(tmp_path):
    n_reads, world = 1001, 2
    port = _free_port()
    mp.spawn(_worker, args=(world, port, n_reads, str(tmp_path)), nprocs=world, join=True)
    counts, locs, off, _, _ = _fake_results(n_reads, 0, world)
    g_off = np.load(tmp_path / "off.npy")
    g_locs = np.load(tmp_path / "locs.npy")
    assert np.array_equal(g_off, off.astype(np.uint32))
    assert np.array_equal(g_locs, locs)


def test_gatherer_equal_shards_world2(tmp_path):
    port = _free_port()
    mp.spawn(_worker, args=(2, port, 2000, str(tmp_path)), nprocs=2, join=True)
    counts, locs, off, _, _ = _fake_results(2000, 0, 2)
    assert np.array_equal(np.load(tmp_path / "off.npy"), off.astype(np.uint32))
    assert np.array_equal(np.load(tmp_path / "locs.npy"), locs)


def test_gather_locations_with_an_empty_rank(tmp_path):
    # 1 read over 2 ranks: rank 0's shard is empty
    port = _free_port()
    mp.spawn(_worker, args=(2, port, 1, str(tmp_path)), nprocs=2, join=True)
    counts, locs, off, _, _ = _fake_results(1, 0, 2)
    assert np.array_equal(np.load(tmp_path / "off.npy"), off.astype(np.uint32))
    assert np.array_equal(np.load(tmp_path / "locs.npy"), locs)
