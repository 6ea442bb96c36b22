"""GPU parity: bucketmap_align verification (Myers bit-vector + traceback) vs the oracle's full DP."""
import numpy as np
import pytest

from common import X, Y, dataset

pytestmark = pytest.mark.gpu

CASES = [
    (dict(read_len=300), dict()),
    (dict(read_len=150), dict()),
    (dict(read_len=300, sub=0.01, indel=0.001), dict()),                      # C5-like error profile
    (dict(read_len=300, sub=0.02, indel=0.02, n_reads=1500), dict(e=0.7, n=0.05)),   # many CIGAR ops (overflow pass)
    (dict(read_len=64, n_reads=1000), dict(r=100)),                           # one 64-bit word exactly
    (dict(read_len=500, n_reads=500), dict()),                                # > 5 words (NW=8 path)
    (dict(rec_split=5, n_bases=1_500_000, read_len=250), dict()),             # short end-of-record buckets clip the window
]


@pytest.mark.parametrize("kw,pk", CASES)
def test_align_matches_oracle(oracle, bm, ctx, kw, pk):
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
    aln, cig = oracle.align(Po, d.packed, d.b_start, d.b_len, d.seq, d.read_off, loff, locs)
    assert len(locs) > 0.5 * (len(d.read_off) - 1)

    ctx.reads_upload(P, d.seq, d.qual, d.read_off)
    ctx.map_device(P)
    ctx.locate_device(P, False)
    goff, glocs = ctx.locate_download()
    assert np.array_equal(goff, loff) and np.array_equal(glocs, locs)
    ctx.align_device(P)
    galn, gcig = ctx.align_download(len(glocs))
    for f in ("score", "begin", "map_qual", "kept", "cigar_len"):
        assert np.array_equal(galn[f], aln[f]), f
    # cigar ops per alignment (offsets may be laid out differently)
    for i in range(len(aln)):
        a = cig[aln["cigar_off"][i]:aln["cigar_off"][i] + aln["cigar_len"][i]]
        g = gcig[galn["cigar_off"][i]:galn["cigar_off"][i] + galn["cigar_len"][i]]
        assert np.array_equal(a, g), i
