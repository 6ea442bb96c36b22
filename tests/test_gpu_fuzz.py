"""GPU: seeded sweep over the run-time parameters (seed lengths, sample counts, fault levels, quality thresholds, read
lengths) -- the table-driven kernels (fused mapper over the coarse k-mer table, locator on the genome k-mer index) and the
table-free kernels against the CPU oracle, bit for bit, on the same inputs."""
import numpy as np
import pytest

from common import X, Y, dataset

pytestmark = pytest.mark.gpu


def _cases(n, seed=20261104):
    rng = np.random.default_rng(seed)
    out = []
    for i in range(n):
        q = int(rng.integers(7, 11))
        k = int(rng.integers(q, min(q + 5, 14) + 1))
        s = int(rng.choice([3, 8, 15, 16, 17, 31, 40, 70]))
        e = float(rng.choice([0.2, 0.4, 0.6, 0.9]))
        p = int(rng.choice([4, 10, 16, 17, 33]))
        b = int(rng.choice([0, 10, 25, 30]))
        d = float(rng.choice([0.3, 0.5, 0.7]))
        read_len = int(rng.choice([60, 150, 300, 500]))
        frac = float(rng.choice([1.0, 0.5]))
        out.append((i, dict(q=q, k=k, s=s, e=e, p=p, b=b, d=d), dict(read_len=read_len, frac=frac, q=q)))
    return out


@pytest.mark.parametrize("idx,pk,kw", _cases(14))
def test_random_parameters_match_oracle(oracle, bm, ctx, idx, pk, kw):
    d = dataset(n_bases=1_200_000, n_reads=600, seed=50 + idx, sub=0.004, indel=0.0005, **kw)
    rng = np.random.default_rng(idx)
    qual = d.qual.copy()
    n = len(d.read_off) - 1
    for i in range(0, n, 2):   # every second read: mixed qualities (the slow path of the k-mer quality test)
        a, e = int(d.read_off[i]), int(d.read_off[i + 1])
        qual[a:e] = (33 + rng.integers(5, 41, e - a)).astype(np.uint8)
    k2i, n_rows = oracle.fracminhash(d.q, X, Y, d.frac)
    rows = oracle.index_build(d.packed, d.b_start, d.b_len, d.NB, d.q, k2i, n_rows)
    Po = oracle.params(d.L, **pk)
    P = bm.default_params(ctx.lib, d.L, **pk)
    segs = oracle.segments(Po, d.read_off)
    coff, cand, stats = oracle.map_segments(Po, d.NB, k2i, rows, d.seq, qual, d.read_off, segs)
    loff, locs = oracle.locate(Po, d.packed, d.b_start, d.b_len, d.seq, qual, d.read_off, segs, coff, cand)
    foff, best = oracle.filter_best(Po, d.read_off, loff, locs)
    ctx.ref_upload(d.packed, d.n_bases, d.b_start, d.b_len, d.NB)
    ctx.index_load(k2i, d.q, rows, d.NB)
    try:
        for tables in (1, 2, 0):                 # 2: coarse table without the full-resolution table
            ctx.index_fine_table(0 if tables == 2 else 1)
            ctx.index_kmer_table(1 if tables else 0)
            ctx.locator_index(1 if tables else 0)
            goff, gcand, gstats = ctx.map_reads(P, d.seq, qual, d.read_off)
            assert np.array_equal(goff, coff) and np.array_equal(gcand, cand), (tables, pk)
            assert np.array_equal(gstats[:4], stats), (tables, pk)
            ctx.locate_device(P, False)
            g2, gl = ctx.locate_download()
            assert np.array_equal(g2, loff) and np.array_equal(gl, locs), (tables, pk)
            ctx.locate_device(P, True)
            g3, gb = ctx.locate_download()
            assert np.array_equal(g3, foff) and np.array_equal(gb, best), (tables, pk)
    finally:
        ctx.index_fine_table(1)
        ctx.index_kmer_table(1)
        ctx.locator_index(1)
