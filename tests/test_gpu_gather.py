"""GPU: the multi-GPU plumbing of the library (bm_comm.cu).  The gather protocol (peer stores into the root's slots, sequence
flags, cuStreamWaitValue32, pack kernel, acknowledgements, two slot sets) runs between two contexts; on a box with one GPU
both contexts sit on that GPU (the peer mapping is then a plain device pointer), with two or more each rank has its own GPU
and the index travels by ncclBroadcast (bm_bcast_index / bm_bcast_ref)."""
import ctypes as C
import os
import threading

import numpy as np
import pytest

import bucketmap_b200 as bm
from common import X, Y, dataset

pytestmark = [pytest.mark.gpu, pytest.mark.timeout(180)]


os.environ.setdefault("BM_GATHER_TIMEOUT_S", "30")   # a protocol bug fails the test instead of hanging the box


def _n_gpus():
    try:
        import torch
        return torch.cuda.device_count()
    except Exception:
        return 1


def _shards(d, world):
    n = len(d.read_off) - 1
    out = []
    for r in range(world):
        lo, hi = n * r // world, n * (r + 1) // world
        a, b = int(d.read_off[lo]), int(d.read_off[hi])
        out.append((d.seq[a:b].copy(), d.qual[a:b].copy(), (d.read_off[lo:hi + 1] - d.read_off[lo]).astype(np.uint64)))
    return out


@pytest.mark.parametrize("world,filter_best", [(2, True), (3, False)])
def test_gather_between_contexts_equals_one_context(oracle, world, filter_best):
    d = dataset(read_len=150, n_reads=3000, n_bases=1_500_000, repeats=20)
    k2i, n_rows = oracle.fracminhash(d.q, X, Y, d.frac)
    rows = oracle.index_build(d.packed, d.b_start, d.b_len, d.NB, d.q, k2i, n_rows)
    ngpu = _n_gpus()
    ctxs = [bm.Context(r % ngpu) for r in range(world)]
    try:
        P = bm.default_params(ctxs[0].lib, d.L)
        for c in ctxs:
            c.ref_upload(d.packed, d.n_bases, d.b_start, d.b_len, d.NB)
            c.index_load(k2i, d.q, rows, d.NB)
        # the whole batch on one context: what the gathered records must equal
        ctxs[0].reads_upload(P, d.seq, d.qual, d.read_off)
        ctxs[0].map_device(P)
        ctxs[0].locate_device(P, filter_best)
        eoff, elocs = ctxs[0].locate_download()
        shards = _shards(d, world)
        mx = max(len(s[2]) - 1 for s in shards)
        hs = [c.gather_open(r, world, 0, mx, 60 * mx) for r, c in enumerate(ctxs)]
        for c in ctxs:
            c.gather_connect(hs)
        n = len(d.read_off) - 1
        for it in range(5):                       # more batches than slot sets: the acknowledgements are exercised
            for r in range(world - 1, -1, -1):    # the root last: with host pointers it waits for the others
                s, q, o = shards[(r + it) % world] if it % 2 else shards[r]
                ctxs[r].reads_upload(P, s, q, o)
                ctxs[r].map_device(P)
                ctxs[r].locate_device(P, filter_best)
                if r == 0 and it == 4:
                    goff = np.empty(n + 1, np.uint32)
                    glocs = np.empty(len(elocs) + 16, bm.LOC_DTYPE)
                    nr, nl = ctxs[0].gather_results(goff, glocs)
                else:
                    ctxs[r].gather_results()
        assert (nr, nl) == (n, len(elocs))
        assert np.array_equal(goff, eoff) and np.array_equal(glocs[:nl], elocs)
        for c in ctxs:
            c.gather_sync()
        o2, l2 = ctxs[0].gather_download()
        assert np.array_equal(o2, eoff) and np.array_equal(l2, elocs)
    finally:
        for c in ctxs:
            c.close()


@pytest.mark.skipif(_n_gpus() < 2, reason="needs two GPUs (NCCL refuses two ranks on one device)")
def test_nccl_broadcast_of_index_and_reference(oracle):
    d = dataset(read_len=150, n_reads=2000, n_bases=1_500_000)
    k2i, n_rows = oracle.fracminhash(d.q, X, Y, d.frac)
    rows = oracle.index_build(d.packed, d.b_start, d.b_len, d.NB, d.q, k2i, n_rows)
    world = 2
    ctxs = [bm.Context(r) for r in range(world)]
    try:
        P = bm.default_params(ctxs[0].lib, d.L)
        ctxs[0].ref_upload(d.packed, d.n_bases, d.b_start, d.b_len, d.NB)
        ctxs[0].index_load(k2i, d.q, rows, d.NB)
        uid = ctxs[0].comm_unique_id()
        errs = []

        def run(r):
            try:
                ctxs[r].comm_init(uid, world, r)
                ctxs[r].bcast_ref(0)
                ctxs[r].bcast_index(0)
            except Exception as e:      # noqa: BLE001
                errs.append(e)
        ts = [threading.Thread(target=run, args=(r,)) for r in range(world)]
        [t.start() for t in ts]
        [t.join() for t in ts]
        assert not errs, errs
        ctxs[1].NB, ctxs[1].n_rows = d.NB, n_rows
        assert np.array_equal(ctxs[1].index_download(), rows)
        res = []
        for c in ctxs:
            c.reads_upload(P, d.seq, d.qual, d.read_off)
            c.map_device(P)
            c.locate_device(P, True)
            res.append(c.locate_download())
        assert np.array_equal(res[0][0], res[1][0]) and np.array_equal(res[0][1], res[1][1])
        assert len(res[0][1]) > 0
    finally:
        for c in ctxs:
            c.close()


@pytest.mark.skipif(_n_gpus() < 2, reason="needs two GPUs (NCCL refuses two ranks on one device)")
def test_sharded_index_build_equals_the_oracle_on_every_rank(oracle):
    """bm_index_build_sharded: each rank indexes its share of the buckets, ncclAllReduce sums the matrices"""
    d = dataset(read_len=150, n_reads=100, n_bases=3_000_000, L=8192, rec_split=3)
    k2i, n_rows = oracle.fracminhash(d.q, X, Y, d.frac)
    rows = oracle.index_build(d.packed, d.b_start, d.b_len, d.NB, d.q, k2i, n_rows)
    world = 2
    ctxs = [bm.Context(r) for r in range(world)]
    try:
        ctxs[0].ref_upload(d.packed, d.n_bases, d.b_start, d.b_len, d.NB)
        uid = ctxs[0].comm_unique_id()
        errs = []

        def run(r):
            try:
                ctxs[r].comm_init(uid, world, r)
                ctxs[r].bcast_ref(0)
                ctxs[r].index_build_sharded(k2i, d.q, n_rows)
            except Exception as e:      # noqa: BLE001
                errs.append(e)
        ts = [threading.Thread(target=run, args=(r,)) for r in range(world)]
        [t.start() for t in ts]
        [t.join() for t in ts]
        assert not errs, errs
        for c in ctxs:
            c.NB, c.n_rows = d.NB, n_rows
            assert np.array_equal(c.index_download(), rows)
    finally:
        for c in ctxs:
            c.close()
