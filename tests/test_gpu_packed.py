"""GPU: the packed read form.  A batch packed on the host (bm_pack_reads) must give the records of the same batch sent
as ASCII (packed on the device by k_pack_reads) and of the CPU oracle; the k-mer quality plane follows (-l, -b)."""
import numpy as np
import pytest

from common import X, Y, dataset

pytestmark = pytest.mark.gpu


def _mixed_quals(d, seed, every=2):
    rng = np.random.default_rng(seed)
    qual = d.qual.copy()
    n = len(d.read_off) - 1
    for i in range(0, n, every):
        a, e = int(d.read_off[i]), int(d.read_off[i + 1])
        qual[a:e] = (33 + rng.integers(5, 41, e - a)).astype(np.uint8)
    return qual


def _setup(oracle, ctx, d):
    k2i, n_rows = oracle.fracminhash(d.q, X, Y, d.frac)
    rows = oracle.index_build(d.packed, d.b_start, d.b_len, d.NB, d.q, k2i, n_rows)
    ctx.ref_upload(d.packed, d.n_bases, d.b_start, d.b_len, d.NB)
    ctx.index_load(k2i, d.q, rows, d.NB)
    return k2i, rows


def _device(ctx, P, filter_best):
    ctx.map_device(P)
    coff, cand, stats = ctx.map_download()
    ctx.locate_device(P, filter_best)
    loff, locs = ctx.locate_download()
    return coff, cand, stats, loff, locs


@pytest.mark.parametrize("kw,pk,mixed", [
    (dict(read_len=300), dict(), False),
    (dict(read_len=300), dict(), True),
    (dict(read_len=150, frac=0.25), dict(b=30), True),
    (dict(read_len=800, n_reads=400), dict(), True),               # long reads: segments start inside chunks
    (dict(read_len=300, sub=0.01, indel=0.001), dict(s=20, e=0.6, k=14, b=10), True),
    (dict(read_len=37, n_reads=500), dict(r=40), True),            # reads of little more than one chunk
])
@pytest.mark.parametrize("tables", [1, 0])
def test_packed_batch_equals_ascii_batch_and_oracle(oracle, bm, ctx, kw, pk, mixed, tables):
    d = dataset(**kw)
    qual = _mixed_quals(d, 11) if mixed else d.qual
    k2i, rows = _setup(oracle, ctx, d)
    Po = oracle.params(d.L, **pk)
    P = bm.default_params(ctx.lib, d.L, **pk)
    segs = oracle.segments(Po, d.read_off)
    ocoff, ocand, ostats = oracle.map_segments(Po, d.NB, k2i, rows, d.seq, qual, d.read_off, segs)
    oloff, olocs = oracle.locate(Po, d.packed, d.b_start, d.b_len, d.seq, qual, d.read_off, segs, ocoff, ocand)
    ctx.index_kmer_table(tables)
    ctx.locator_index(tables)
    try:
        ctx.reads_upload(P, d.seq, qual, d.read_off)
        a = _device(ctx, P, False)
        planes, qk, pl_off, low = bm.pack_reads(ctx.lib, d.seq, qual, d.read_off, P.k, P.avg_base_quality)
        ctx.reads_upload_packed(P, planes, qk if low else None, pl_off, d.read_off)
        b = _device(ctx, P, False)
        for i, (x, y) in enumerate(zip(a, b)):
            assert np.array_equal(x[:4], y[:4]) if i == 2 else np.array_equal(x, y)   # stats[4] (bytes loaded) differs
        assert np.array_equal(a[0], ocoff) and np.array_equal(a[1], ocand) and np.array_equal(a[2][:4], ostats)
        assert np.array_equal(a[3], oloff) and np.array_equal(a[4], olocs)
        # the quality plane always travels when given, even if all-pass
        if not low:
            ctx.reads_upload_packed(P, planes, qk, pl_off, d.read_off)
            c = _device(ctx, P, False)
            for i, (x, y) in enumerate(zip(a, c)):
                assert np.array_equal(x[:4], y[:4]) if i == 2 else np.array_equal(x, y)
    finally:
        ctx.index_kmer_table(1)
        ctx.locator_index(1)


def test_map_locate_packed_equals_map_locate(oracle, bm, ctx):
    """host-buffer entry points, above and below the pipelining threshold (32768 reads)"""
    for n_reads in (2000, 70_000):
        d = dataset(read_len=150, n_reads=n_reads, n_bases=3_000_000)
        qual = _mixed_quals(d, 3, every=5)
        _setup(oracle, ctx, d)
        P = bm.default_params(ctx.lib, d.L)
        off_a, locs_a, st_a = ctx.map_locate(P, d.seq, qual, d.read_off, True)
        planes, qk, pl_off, low = bm.pack_reads(ctx.lib, d.seq, qual, d.read_off, P.k, P.avg_base_quality)
        assert low > 0
        n = len(d.read_off) - 1
        off_b = np.empty(n + 1, np.uint32)
        locs_b = np.empty(len(locs_a) + 16, bm.LOC_DTYPE)
        nl, st_b = ctx.map_locate_packed_into(P, planes, qk, pl_off, d.read_off, off_b, locs_b, True)
        assert nl == len(locs_a)
        assert np.array_equal(off_a, off_b) and np.array_equal(locs_a, locs_b[:nl])
        assert np.array_equal(st_a[:4], st_b[:4])
        # a slice of the batch through the packed entry point (pl_off[0] != 0)
        lo, hi = n // 3, n // 3 + 40_000 if n_reads > 40_000 else n // 2
        off_c = np.empty(hi - lo + 1, np.uint32)
        locs_c = np.empty(len(locs_a) + 16, bm.LOC_DTYPE)
        nc, _ = ctx.map_locate_packed_into(P, planes, qk, pl_off[lo:hi + 1], d.read_off[lo:hi + 1], off_c, locs_c, True)
        assert np.array_equal(np.diff(off_c), np.diff(off_a[lo:hi + 1]))
        assert np.array_equal(locs_c[:nc], locs_a[off_a[lo]:off_a[hi]])


def test_quality_plane_follows_the_parameters(oracle, bm, ctx):
    """the plane is per (-l, -b): a resident ASCII batch is re-packed when the next call brings other values; a packed
    batch refuses a higher -b than it was packed for"""
    d = dataset(read_len=300)
    qual = _mixed_quals(d, 7)
    k2i, rows = _setup(oracle, ctx, d)
    P1 = bm.default_params(ctx.lib, d.L)
    P2 = bm.default_params(ctx.lib, d.L, b=12, k=13)
    ctx.reads_upload(P1, d.seq, qual, d.read_off)
    for P, kw in ((P1, dict()), (P2, dict(b=12, k=13)), (P1, dict())):
        Po = oracle.params(d.L, **kw)
        segs = oracle.segments(Po, d.read_off)
        ocoff, ocand, _ = oracle.map_segments(Po, d.NB, k2i, rows, d.seq, qual, d.read_off, segs)
        oloff, olocs = oracle.locate(Po, d.packed, d.b_start, d.b_len, d.seq, qual, d.read_off, segs, ocoff, ocand)
        coff, cand, _, loff, locs = _device(ctx, P, False)
        assert np.array_equal(coff, ocoff) and np.array_equal(cand, ocand)
        assert np.array_equal(loff, oloff) and np.array_equal(locs, olocs)
    planes, qk, pl_off, low = bm.pack_reads(ctx.lib, d.seq, d.qual, d.read_off, P1.k, P1.avg_base_quality)
    assert low == 0
    ctx.reads_upload_packed(P1, planes, None, pl_off, d.read_off)
    ctx.map_device(P2)                                       # lower -b: "all windows pass" still holds
    P3 = bm.default_params(ctx.lib, d.L, b=40)
    with pytest.raises(bm.BucketMapError):
        ctx.map_device(P3)


def test_pipelined_packed_call_with_a_long_read_in_a_late_chunk(oracle, bm, ctx):
    """bm_map_locate_packed scans the read offsets chunk by chunk assuming one segment per read; a read longer than 2r found
    in a later chunk makes the call start over with the general segment tables.  Same records as the ASCII call."""
    d = dataset(read_len=150, n_reads=70_000, n_bases=3_000_000)
    _setup(oracle, ctx, d)
    P = bm.default_params(ctx.lib, d.L)
    # reads 60000..60004 become one 750-base read (> 2r = 600): five segments
    read_off = np.delete(d.read_off, [60001, 60002, 60003, 60004])
    n = len(read_off) - 1
    assert int(read_off[60001] - read_off[60000]) == 750
    off_a, locs_a, st_a = ctx.map_locate(P, d.seq, d.qual, read_off, True)
    planes, qk, pl_off, low = bm.pack_reads(ctx.lib, d.seq, d.qual, read_off, P.k, P.avg_base_quality)
    off_b = np.empty(n + 1, np.uint32)
    locs_b = np.empty(len(locs_a) + 16, bm.LOC_DTYPE)
    nl, st_b = ctx.map_locate_packed_into(P, planes, qk if low else None, pl_off, read_off, off_b, locs_b, True)
    assert nl == len(locs_a)
    assert np.array_equal(off_a, off_b) and np.array_equal(locs_a, locs_b[:nl])
    assert np.array_equal(st_a[:4], st_b[:4])
    assert len(ctx.segments()[0]) == n + 4


def test_pipelined_packed_call_rejects_a_bad_layout_in_a_late_chunk(oracle, bm, ctx):
    d = dataset(read_len=150, n_reads=70_000, n_bases=3_000_000)
    _setup(oracle, ctx, d)
    P = bm.default_params(ctx.lib, d.L)
    planes, qk, pl_off, low = bm.pack_reads(ctx.lib, d.seq, d.qual, d.read_off, P.k, P.avg_base_quality)
    n = len(d.read_off) - 1
    bad = pl_off.copy()
    bad[65_000] += 1
    off_b = np.empty(n + 1, np.uint32)
    locs_b = np.empty(4 * n, bm.LOC_DTYPE)
    with pytest.raises(bm.BucketMapError, match="pl_off"):
        ctx.map_locate_packed_into(P, planes, None, bad, d.read_off, off_b, locs_b, True)
    # the context is usable afterwards
    nl, _ = ctx.map_locate_packed_into(P, planes, None, pl_off, d.read_off, off_b, locs_b, True)
    assert nl > 0.9 * n


def test_lane_context_gives_the_same_records_from_a_second_thread(oracle, bm, ctx):
    """bm_ctx_create_lane: a context viewing the parent's index; two host threads keep two batches in flight"""
    import threading
    d = dataset(read_len=150, n_reads=70_000, n_bases=3_000_000)
    _setup(oracle, ctx, d)
    P = bm.default_params(ctx.lib, d.L)
    planes, qk, pl_off, low = bm.pack_reads(ctx.lib, d.seq, d.qual, d.read_off, P.k, P.avg_base_quality)
    n = len(d.read_off) - 1
    off_a = np.empty(n + 1, np.uint32)
    locs_a = np.empty(4 * n, bm.LOC_DTYPE)
    na, _ = ctx.map_locate_packed_into(P, planes, None, pl_off, d.read_off, off_a, locs_a, True)
    want_off, want = off_a.copy(), locs_a[:na].copy()
    lane = ctx.create_lane(P)
    try:
        off_b = np.empty(n + 1, np.uint32)
        locs_b = np.empty(4 * n, bm.LOC_DTYPE)
        errs = []

        def drive(c, oo, ol, lo, hi):
            try:
                for _ in range(4):
                    m, _ = c.map_locate_packed_into(P, planes, None, pl_off[lo:hi + 1], d.read_off[lo:hi + 1], oo, ol, True)
                    assert np.array_equal(np.diff(oo[:hi - lo + 1]), np.diff(want_off[lo:hi + 1]))
                    assert np.array_equal(ol[:m], want[want_off[lo]:want_off[hi]])
            except Exception as e:                       # noqa: BLE001 -- reported by the main thread
                errs.append(e)

        th = [threading.Thread(target=drive, args=(ctx, off_a, locs_a, 0, n)),
              threading.Thread(target=drive, args=(lane, off_b, locs_b, 5_000, 65_000))]
        for t in th:
            t.start()
        for t in th:
            t.join()
        assert not errs, errs
        with pytest.raises(bm.BucketMapError, match="lane"):
            lane.create_lane(P)
    finally:
        lane.close()
