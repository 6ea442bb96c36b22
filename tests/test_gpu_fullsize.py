"""GPU: BASELINE.json's headline configuration at full size (1.7 Gbp reference, L=65536, -f 1, 1 M x 300 bp reads) --
too large for the CPU oracle end to end, so checked through size-independent properties plus a bit-exact oracle
comparison on a prefix of the full batch (same index, same genome)."""
import numpy as np
import pytest

import bucketmap_b200 as bm
from bucketmap_b200 import datagen, hostlogic

pytestmark = pytest.mark.gpu

N_BASES, L, Q = 1_700_000_000, 65536, 9
N_READS = 1_000_000
X, Y = 40503, 7919


class Full:
    pass


@pytest.fixture(scope="module")
def full():
    f = Full()
    f.ctx = bm.Context(0)
    f.P = bm.default_params(f.ctx.lib, L)
    f.packed = datagen.random_genome_packed(N_BASES, 3)
    f.NB, f.b_start, f.b_len, _ = datagen.bucketize(np.array([N_BASES], np.uint64), L, 300)
    f.k2i, f.n_rows = hostlogic.fracminhash(Q, X, Y, 1.0)
    f.ctx.ref_upload(f.packed, N_BASES, f.b_start, f.b_len, f.NB)
    f.ctx.index_build(f.k2i, Q, f.n_rows, download=False)
    f.seq, f.qual, f.off, f.start, f.rev = datagen.simulate_reads(f.packed, N_BASES, N_READS, 300, 0.002, 0.00025,
                                                                   0.00025, 4)
    yield f
    f.ctx.close()


def _run(ctx, P, seq, qual, off, filter_best=True):
    ctx.reads_upload(P, seq, qual, off)
    ctx.map_device(P)
    coff, cand, stats = ctx.map_download()
    ctx.locate_device(P, filter_best)
    loff, locs = ctx.locate_download()
    return coff, cand, stats, loff, locs


def test_full_index_columns_hold_the_distinct_qgrams_of_their_bucket(full):
    """checksum of the 850 MB matrix: column b has one bit per distinct 9-mer of bucket b (every 9-mer is a row at
    -f 1), and the total number of set bits is the sum over buckets"""
    assert full.NB == 25940 and len(full.b_start) == 25940
    rows = full.ctx.index_download()
    assert rows.shape == (4 ** Q, (full.NB + 7) >> 3)
    for b in (0, 1, 777, 12970, 25938, 25939):
        r = datagen.unpack(full.packed, int(full.b_start[b]), int(full.b_len[b])).astype(np.int64)
        h = np.zeros(len(r) - Q + 1, np.int64)
        for i in range(Q):
            h = h * 4 + r[i:len(r) - Q + 1 + i]
        distinct = np.unique(h)
        col = (rows[:, b >> 3] >> (b & 7)) & 1
        assert int(col.sum()) == len(distinct)
        assert np.array_equal(np.nonzero(col)[0], distinct)           # row id == 9-mer code at -f 1
    # round trip through the .qgram layout: load what was downloaded, download again
    z, valid = full.ctx.index_zeros(0.5)
    ones = np.unpackbits(rows[:4096], axis=1, bitorder="little").sum(axis=1)
    assert np.array_equal(z[:4096], full.NB - ones)
    del rows


def test_full_batch_equals_its_shards_and_the_oracle_prefix(full, oracle):
    coff, cand, stats, loff, locs = _run(full.ctx, full.P, full.seq, full.qual, full.off)
    assert int(stats[3]) == 120 * N_READS                              # R = 2*S*(k-q+1) rows per read at -f 1
    # idempotence
    coff2, cand2, _, loff2, locs2 = _run(full.ctx, full.P, full.seq, full.qual, full.off)
    assert np.array_equal(coff, coff2) and np.array_equal(cand, cand2)
    assert np.array_equal(loff, loff2) and np.array_equal(locs, locs2)
    # sharding invariance (what the multi-GPU path relies on): 4 contiguous shards, concatenated
    parts_c, parts_l, cnt_c, cnt_l = [], [], [], []
    for s in range(4):
        lo, hi = N_READS * s // 4, N_READS * (s + 1) // 4
        o = (full.off[lo:hi + 1] - full.off[lo]).astype(np.uint64)
        a, b = int(full.off[lo]), int(full.off[hi])
        c_off, c, _, l_off, l = _run(full.ctx, full.P, full.seq[a:b], full.qual[a:b], o)
        parts_c.append(c); parts_l.append(l)
        cnt_c.append(np.diff(c_off)); cnt_l.append(np.diff(l_off))
    assert np.array_equal(np.concatenate(parts_c), cand)
    assert np.array_equal(np.concatenate(cnt_c), np.diff(coff))
    assert np.array_equal(np.concatenate(parts_l), locs)
    assert np.array_equal(np.concatenate(cnt_l), np.diff(loff))
    # accuracy against the simulator's ground truth (README: 97 % correct positions on the real genome)
    first = loff[:-1]
    has = np.diff(loff) > 0
    lb = locs["bucket"][first[has]].astype(np.int64) * L + locs["offset"][first[has]]
    ok = np.abs(lb - full.start[has]) <= 8
    strand_ok = locs["is_original"][first[has]] == (~full.rev[has]).astype(np.uint32)
    assert has.mean() > 0.99 and (ok & strand_ok).mean() > 0.99
    # bit-exact against the CPU oracle on a prefix of the full batch (full-size index and genome)
    npre = 3000
    rows = full.ctx.index_download()
    Po = oracle.params(L)
    sub_off = full.off[:npre + 1].copy()
    segs = oracle.segments(Po, sub_off)
    ocoff, ocand, _ = oracle.map_segments(Po, full.NB, full.k2i, rows, full.seq, full.qual, sub_off, segs)
    assert np.array_equal(coff[:2 * npre + 1], ocoff) and np.array_equal(cand[:ocoff[-1]], ocand)
    oloff, olocs = oracle.locate(Po, full.packed, full.b_start, full.b_len, full.seq, full.qual, sub_off, segs, ocoff, ocand)
    ofoff, obest = oracle.filter_best(Po, sub_off, oloff, olocs)
    assert np.array_equal(loff[:npre + 1], ofoff) and np.array_equal(locs[:ofoff[-1]], obest)


def test_full_error_free_reads_land_on_their_origin_and_align_perfectly(full):
    n = 200_000
    seq, qual, off, start, rev = datagen.simulate_reads(full.packed, N_BASES, n, 300, 0.0, 0.0, 0.0, 77)
    ctx, P = full.ctx, full.P
    ctx.reads_upload(P, seq, qual, off)
    ctx.map_device(P)
    ctx.locate_device(P, False)
    loff, locs = ctx.locate_download()
    ctx.align_device(P)
    aln, cig = ctx.align_download(len(locs))
    cnt = np.diff(loff)
    # every read is found, except those starting exactly at a bucket offset 0 (dropped by `offset > 0`,
    # bucket_locator.h:674,686) -- which can still be found in the previous bucket's overlap
    assert (cnt > 0).mean() > 0.9999
    read_of = np.repeat(np.arange(n), cnt)
    pos = locs["bucket"].astype(np.int64) * L + locs["offset"]
    exact = (pos == start[read_of]) & (locs["is_original"] == (~rev[read_of]).astype(np.uint32))
    # a read hanging over the end of a bucket is still located there (its first samples vote), but its text window is
    # clipped at the bucket end (bucket_locator.h:550): only whole-read windows must verify perfectly
    whole = locs["offset"].astype(np.int64) + 300 <= full.b_len[locs["bucket"]].astype(np.int64)
    found = np.zeros(n, bool)
    np.logical_or.at(found, read_of[exact], True)
    assert found[cnt > 0].all()
    exact &= whole
    assert exact.sum() >= (cnt > 0).sum()
    # an exact location verifies with edit distance 0, begin 0 and a single 300M op
    assert (aln["score"][exact] == 0).all()
    # forward strand: the read starts at column 0 of its text window; reverse strand: the window (read length + 1 +
    # (size_t)(0.02 * 300) = 307 columns, clipped at the bucket end) is reverse-complemented, so the read sits at its end
    fwd = exact & (locs["is_original"] == 1)
    rv = exact & (locs["is_original"] == 0)
    assert (aln["begin"][fwd] == 0).all()
    width = np.minimum(307, full.b_len[locs["bucket"][rv]].astype(np.int64) - locs["offset"][rv])
    assert np.array_equal(aln["begin"][rv].astype(np.int64), width - 300)
    assert (aln["kept"][exact] == 1).all() and (aln["map_qual"][exact] == 60).all()
    assert (aln["cigar_len"][exact] == 1).all()
    assert (cig[aln["cigar_off"][exact]] == ((300 << 4) | 0)).all()
    # scores are never positive, and map_qual is 60 + score in 32-bit arithmetic
    assert (aln["score"] <= 0).all()
    assert np.array_equal(aln["map_qual"], (60 + aln["score"].astype(np.int64)).astype(np.uint32))
