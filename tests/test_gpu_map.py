"""GPU parity: index build / load / row zeros / mapper candidate lists vs the CPU oracle (bit-exact)."""
import numpy as np
import pytest

from common import X, Y, dataset

pytestmark = pytest.mark.gpu


def _index(oracle, d, frac=1.0):
    k2i, n_rows = oracle.fracminhash(d.q, X, Y, frac)
    rows = oracle.index_build(d.packed, d.b_start, d.b_len, d.NB, d.q, k2i, n_rows)
    return k2i, n_rows, rows


@pytest.mark.parametrize("kw", [dict(), dict(rec_split=5, n_bases=1_500_000), dict(frac=0.25), dict(L=16384, n_bases=3_000_000)])
def test_index_build_matches_oracle(oracle, bm, ctx, kw):
    d = dataset(**kw)
    k2i, n_rows, rows = _index(oracle, d, d.frac)
    ctx.ref_upload(d.packed, d.n_bases, d.b_start, d.b_len, d.NB)
    got = ctx.index_build(k2i, d.q, n_rows)
    assert got.shape == rows.shape
    assert np.array_equal(got, rows)
    z, v = ctx.index_zeros(0.5)
    zo, vo = oracle.zeros(rows, d.NB, 0.5)
    assert np.array_equal(z, zo) and v == vo


def test_index_load_roundtrip(oracle, bm, ctx):
    d = dataset()
    k2i, n_rows, rows = _index(oracle, d)
    ctx.ref_upload(d.packed, d.n_bases, d.b_start, d.b_len, d.NB)
    ctx.index_load(k2i, d.q, rows, d.NB)
    assert np.array_equal(ctx.index_download(), rows)


@pytest.mark.parametrize("kw,pk", [
    (dict(), dict()),
    (dict(read_len=300), dict()),
    (dict(frac=0.25, read_len=300), dict()),
    (dict(rec_split=5, n_bases=1_500_000), dict()),
    (dict(L=16384, n_bases=3_000_000, read_len=300, n_reads=3000), dict()),
    (dict(read_len=300, repeats=40), dict()),
    (dict(read_len=300, sub=0.01, indel=0.001), dict(s=20, e=0.6, k=14, b=10)),
    (dict(read_len=800, n_reads=500), dict()),           # long reads: 5 segments each
])
# -1 / -2: table on, every list / lists with > 1 passing group handed to K2; -3: coarse table on, full-resolution table off
# (passing groups recounted from the .qgram words)
@pytest.mark.parametrize("ktable", [1, 0, 8, 32, -1, -2, -3])
def test_map_matches_oracle(oracle, bm, ctx, kw, pk, ktable):
    d = dataset(**kw)
    handover = {-1: 0, -2: 1}.get(ktable, -1)
    fine = 0 if ktable == -3 else 1
    ktable = 16 if ktable < 0 else ktable     # the hand-over / fine-table variants run on a forced table (1 = the cost model's choice)
    ctx.index_fine_table(fine)
    ctx.index_kmer_table(ktable)
    ctx.map_handover(handover)
    k2i, n_rows, rows = _index(oracle, d, d.frac)
    Po = oracle.params(d.L, **pk)
    P = bm.default_params(ctx.lib, d.L, **pk)
    ctx.ref_upload(d.packed, d.n_bases, d.b_start, d.b_len, d.NB)
    ctx.index_load(k2i, d.q, rows, d.NB)
    segs = oracle.segments(Po, d.read_off)
    coff, cand, stats = oracle.map_segments(Po, d.NB, k2i, rows, d.seq, d.qual, d.read_off, segs)
    goff, gcand, gstats = ctx.map_reads(P, d.seq, d.qual, d.read_off)
    gs = ctx.segments()
    for a, b in zip(segs, gs):
        assert np.array_equal(a, b)
    assert np.array_equal(goff, coff)
    assert np.array_equal(gcand, cand)
    assert np.array_equal(gstats[:4], stats)
    assert stats[0] > 0.9 * len(segs[0]) or kw.get("read_len") == 800
    tk, tbytes = ctx.index_kmer_table_info()
    fk, fbytes = ctx.index_fine_table_info()
    chosen, density = ctx.index_kmer_table_choice()
    if ktable == 1:                               # mode 1: 8 / 16 / 32 buckets per coarse bit, or no table when the gather is cheaper
        assert chosen in (0, 8, 16, 32) and 0.0 < density <= 1.0
        if kw.get("frac", 1.0) < 0.5:
            # FracMinHash-sparse index: every coarse bit would be set and the gather is cheaper (C2 at -f 0.25, bench.py) --
            # unless the matrix is small enough to live in L2, as every matrix of this file is: then the fused kernel is kept
            assert chosen == 16
    else:
        assert chosen == ktable
    have = chosen != 0
    assert (fk, fbytes) == ((P.k, 4 ** P.k * ((((d.NB + 7) >> 3) + 63) & ~63)) if (have and fine) else (0, 0))
    ctx.index_fine_table(1)
    ctx.index_kmer_table(1)   # session-wide context: back to the default for the tests that follow
    ctx.map_handover(-1)
    wp = (((d.NB + 7) >> 3) + 63) & ~63                       # padded row bytes
    wcs = ((wp * 8 // max(chosen, 1) + 31) // 32 + 15) & ~15   # coarse row stride in 32-bit words
    assert (tk, tbytes) == ((P.k, (4 ** P.k + 1) * wcs * 4) if have else (0, 0))   # + the all-zero padding row
