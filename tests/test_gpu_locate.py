"""GPU parity: locator (sampling, in-bucket offset vote, per-read order, best-location merge) vs the CPU oracle."""
import numpy as np
import pytest

from common import X, Y, dataset

pytestmark = pytest.mark.gpu

CASES = [
    (dict(), dict()),
    (dict(read_len=300), dict()),
    (dict(frac=0.25, read_len=300), dict()),
    (dict(rec_split=5, n_bases=1_500_000), dict()),
    (dict(L=16384, n_bases=3_000_000, read_len=300, n_reads=3000), dict()),
    (dict(read_len=300, repeats=40), dict()),
    (dict(read_len=300, sub=0.01, indel=0.001), dict(s=20, e=0.6, k=14, b=10)),
    (dict(read_len=800, n_reads=500), dict()),            # long reads: 5 segments, merged by _filter_best_locations
    (dict(read_len=300, tandem=20000, n_reads=800), dict()),   # low complexity: hit spill + dense vote table
    (dict(read_len=150, n_bases=300_000, n_reads=6000), dict()),  # many candidates per bucket (several chunks)
]


# (coarse k-mer table + fused mapper kernel, genome k-mer position index) on / off: every combination is a CUDA path
@pytest.mark.parametrize("modes", [(1, 1), (0, 0), (1, 0), (0, 1)])
@pytest.mark.parametrize("kw,pk", CASES)
def test_locate_matches_oracle(oracle, bm, ctx, kw, pk, modes):
    ctx.index_kmer_table(modes[0])
    ctx.locator_index(modes[1])
    try:
        _locate_case(oracle, bm, ctx, kw, pk)
    finally:
        ctx.index_kmer_table(1)
        ctx.locator_index(1)


def _locate_case(oracle, bm, ctx, kw, pk):
    d = dataset(**kw)
    k2i, n_rows = oracle.fracminhash(d.q, X, Y, d.frac)
    rows = oracle.index_build(d.packed, d.b_start, d.b_len, d.NB, d.q, k2i, n_rows)
    Po = oracle.params(d.L, **pk)
    P = bm.default_params(ctx.lib, d.L, **pk)
    ctx.ref_upload(d.packed, d.n_bases, d.b_start, d.b_len, d.NB)
    ctx.index_load(k2i, d.q, rows, d.NB)
    segs = oracle.segments(Po, d.read_off)
    coff, cand, _ = oracle.map_segments(Po, d.NB, k2i, rows, d.seq, d.qual, d.read_off, segs)
    loff, locs = oracle.locate(Po, d.packed, d.b_start, d.b_len, d.seq, d.qual, d.read_off, segs, coff, cand)
    foff, best = oracle.filter_best(Po, d.read_off, loff, locs)

    ctx.reads_upload(P, d.seq, d.qual, d.read_off)
    ctx.map_device(P)
    ctx.locate_device(P, False)
    goff, glocs = ctx.locate_download()
    assert np.array_equal(goff, loff)
    assert np.array_equal(glocs, locs)
    assert len(locs) > 0
    ctx.locate_device(P, True)
    goff, glocs = ctx.locate_download()
    assert np.array_equal(goff, foff)
    assert np.array_equal(glocs, best)
    # host-buffer entry point gives the same
    hoff, hlocs, _ = ctx.map_locate(P, d.seq, d.qual, d.read_off, True)
    assert np.array_equal(hoff, foff) and np.array_equal(hlocs, best)


def test_locator_samples_match_oracle(oracle, bm, ctx):
    """low qualities: the quality filter and the all-k-mers fallback of _prepare_read_query"""
    d = dataset(read_len=300)
    rng = np.random.default_rng(9)
    qual = d.qual.copy()
    n = len(d.read_off) - 1
    for i in range(0, n, 3):   # every third read: mixed qualities; every ninth: all low
        a, b = int(d.read_off[i]), int(d.read_off[i + 1])
        qual[a:b] = (33 + rng.integers(0, 41, b - a)).astype(np.uint8) if i % 9 else 33 + 5
    k2i, n_rows = oracle.fracminhash(d.q, X, Y, 1.0)
    rows = oracle.index_build(d.packed, d.b_start, d.b_len, d.NB, d.q, k2i, n_rows)
    Po = oracle.params(d.L)
    P = bm.default_params(ctx.lib, d.L)
    ctx.ref_upload(d.packed, d.n_bases, d.b_start, d.b_len, d.NB)
    ctx.index_load(k2i, d.q, rows, d.NB)
    segs = oracle.segments(Po, d.read_off)
    coff, cand, stats = oracle.map_segments(Po, d.NB, k2i, rows, d.seq, qual, d.read_off, segs)
    goff, gcand, gstats = ctx.map_reads(P, d.seq, qual, d.read_off)
    assert np.array_equal(goff, coff) and np.array_equal(gcand, cand) and np.array_equal(gstats[:4], stats)
    loff, locs = oracle.locate(Po, d.packed, d.b_start, d.b_len, d.seq, qual, d.read_off, segs, coff, cand)
    ctx.locate_device(P, False)
    g2, gl = ctx.locate_download()
    assert np.array_equal(g2, loff) and np.array_equal(gl, locs)


@pytest.mark.parametrize("lanes", [0xFFFFFFFF, 65536])   # one pipeline lane / the batch cut in two lanes
@pytest.mark.parametrize("filter_best", [True, False])
def test_pipelined_map_locate_matches_oracle_and_serial_path(oracle, bm, ctx, filter_best, lanes):
    """bm_map_locate cuts batches of >= 32768 reads into chunks (and large ones over two lanes): same records, same order"""
    ctx.pipeline_lanes(lanes)
    try:
        _pipelined_case(oracle, bm, ctx, filter_best)
    finally:
        ctx.pipeline_lanes(0)


def _pipelined_case(oracle, bm, ctx, filter_best):
    d = dataset(n_bases=3_000_000, n_reads=70_001, read_len=150, repeats=20)
    k2i, n_rows = oracle.fracminhash(d.q, X, Y, d.frac)
    rows = oracle.index_build(d.packed, d.b_start, d.b_len, d.NB, d.q, k2i, n_rows)
    Po = oracle.params(d.L)
    P = bm.default_params(ctx.lib, d.L)
    ctx.ref_upload(d.packed, d.n_bases, d.b_start, d.b_len, d.NB)
    ctx.index_load(k2i, d.q, rows, d.NB)
    segs = oracle.segments(Po, d.read_off)
    coff, cand, ostats = oracle.map_segments(Po, d.NB, k2i, rows, d.seq, d.qual, d.read_off, segs)
    loff, locs = oracle.locate(Po, d.packed, d.b_start, d.b_len, d.seq, d.qual, d.read_off, segs, coff, cand)
    if filter_best:
        loff, locs = oracle.filter_best(Po, d.read_off, loff, locs)
    n = len(d.read_off) - 1
    out_off = np.zeros(n + 1, np.uint32)
    out_locs = np.zeros(len(cand) + 16, bm.LOC_DTYPE)
    n_locs, stats = ctx.map_locate_into(P, d.seq, d.qual, d.read_off, out_off, out_locs, filter_best)
    assert n_locs == len(locs)
    assert np.array_equal(out_off, loff)
    assert np.array_equal(out_locs[:n_locs], locs)
    assert np.array_equal(stats[:4], ostats)
    # serial device path
    ctx.reads_upload(P, d.seq, d.qual, d.read_off)
    ctx.map_device(P)
    ctx.locate_device(P, filter_best)
    soff, slocs = ctx.locate_download()
    assert np.array_equal(soff, loff) and np.array_equal(slocs, locs)
    # a too-small output buffer is an error, not a truncation
    with pytest.raises(bm.BucketMapError):
        ctx.map_locate_into(P, d.seq, d.qual, d.read_off, out_off, out_locs[:100], filter_best)
