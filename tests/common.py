"""Seeded datasets shared by the tests (oracle side computed once per session and cached)."""
import functools

import numpy as np

import bucketmap_b200 as bm
from bucketmap_b200 import datagen

X, Y = 40503, 7919   # fixed FracMinHash hash parameters (the reference draws them from rand())


class Dataset:
    pass


@functools.lru_cache(maxsize=None)
def dataset(n_bases=2_000_000, L=65536, n_reads=2000, read_len=150, frac=1.0, seed=1, sub=0.002, indel=0.00025,
            rec_split=1, repeats=0, r=300, q=9, tandem=0):
    """genome + buckets + FracMinHash table + simulated reads.  rec_split>1 cuts the genome into several FASTA
    records (so some records end in short / skipped buckets)."""
    d = Dataset()
    d.n_bases, d.L, d.r, d.q, d.frac = n_bases, L, r, q, frac
    d.packed = datagen.random_genome_packed(n_bases, seed)
    if repeats:
        d.packed = datagen.add_repeats(d.packed, n_bases, 3000, repeats, seed + 100)
    if tandem:
        # low-complexity stretch: a short unit repeated over `tandem` bases, and a share of the reads drawn from it
        # (k-mers with thousands of occurrences per bucket: the locator's spill paths)
        rng = np.random.default_rng(seed + 55)
        ranks = datagen.unpack(d.packed, 0, n_bases).astype(np.uint8)
        unit = rng.integers(0, 4, 7).astype(np.uint8)
        t0 = n_bases // 3
        ranks[t0:t0 + tandem] = np.resize(unit, tandem)
        d.packed = datagen.pack(ranks)
        d.tandem_start = t0
    if rec_split == 1:
        d.rec_len = np.array([n_bases], np.uint64)
    else:
        rng = np.random.default_rng(seed + 7)
        cuts = np.sort(rng.integers(1, n_bases, rec_split - 1))
        d.rec_len = np.diff(np.concatenate([[0], cuts, [n_bases]])).astype(np.uint64)
    d.names = [f"chr{i} synthetic record {i}" for i in range(len(d.rec_len))]
    d.NB, d.b_start, d.b_len, d.b_rec = datagen.bucketize(d.rec_len, L, r)
    d.seq, d.qual, d.read_off, d.truth_start, d.truth_rev = datagen.simulate_reads(
        d.packed, n_bases, n_reads, read_len, sub, indel, indel, seed + 1,
        hot=(d.tandem_start - read_len, d.tandem_start + tandem) if tandem else None)
    return d


def lists_equal(off_a, a, off_b, b):
    return np.array_equal(off_a, off_b) and np.array_equal(a, b)
