"""CPU: host-side integer logic of the product (bucket-map_b200/hostlogic.py, datagen.bucketize) against the oracle, and
the oracle's own corner cases listed in SURVEY.md 8(a)-notes that need no GPU."""
import numpy as np
import pytest

from bucketmap_b200 import datagen, hostlogic


@pytest.mark.parametrize("frac", [1.0, 0.25, 0.5, 0.1, 0.0])
@pytest.mark.parametrize("xy", [(40503, 7919), (1, 0), (116730, 116730), (99991, 5)])
def test_fracminhash_matches_oracle(oracle, frac, xy):
    k2i, n = hostlogic.fracminhash(9, xy[0], xy[1], frac)
    ok2i, on = oracle.fracminhash(9, xy[0], xy[1], frac)
    assert n == on and np.array_equal(k2i, ok2i)
    if frac == 1.0:
        assert n == 4 ** 9                       # threshold 10000 >= every hash value
    if frac == 0.25:
        assert 0.24 < n / 4 ** 9 < 0.27          # bucketmap_fracMinHash_map.log: 25.7 % kept


def test_fracminhash_threshold_is_float_truncated(oracle):
    # (unsigned)(10000 * 0.25f) = 2500 and the comparison is <=, so 2501 of 10000 hash values are kept
    k2i, n = oracle.fracminhash(6, 7, 3, 0.25)
    h = (7 * np.arange(4 ** 6, dtype=np.uint64) + 3) % 116731 % 10000
    assert n == int((h <= 2500).sum())


@pytest.mark.parametrize("rec_len,L,r", [
    ([4_600_000], 65536, 300),
    ([100_000, 299, 300, 301, 65536, 65537, 65836, 65837, 131072 + 300, 131072 + 301], 65536, 300),
    ([1, 16384, 16385, 50_000], 16384, 150),
    ([1_000_000], 262144, 300),
])
def test_bucketize_matches_oracle(oracle, rec_len, L, r):
    rl = np.array(rec_len, np.uint64)
    NB, bs, bl, br = datagen.bucketize(rl, L, r)
    oNB, obs, obl, obr = oracle.bucketize(rl, L, r)
    assert NB == oNB == hostlogic.num_buckets(rl, L)
    assert np.array_equal(bs, obs) and np.array_equal(bl, obl) and np.array_equal(br, obr)
    assert len(bs) <= NB                        # skipped short tails leave phantom bucket ids (SURVEY 8a note 4)
    assert (bl > r).all() and (bl <= L + r).all()


def test_sampler_matches_reference_formula(oracle):
    # utils.h:160-178: floor(i * (ub+1)/(n-1)) for i < n-1, then ub
    for n, ub in [(15, 288), (15, 2), (10, 288), (10, 5), (5, 499), (20, 138), (1, 7), (15, 14)]:
        s = oracle.sample(n, ub)
        delta = 0.0 if n == 1 else (ub + 1) / (n - 1)
        want = [int(np.floor(i * delta)) for i in range(n - 1)] + [ub]
        assert list(s) == want
        assert max(s) <= ub + 0 or n == 1
    # fewer candidates than samples: duplicates appear (SURVEY 8a note 2)
    assert len(set(oracle.sample(15, 3))) < 15


def test_segments_long_reads(oracle):
    P = oracle.params(65536)
    off = np.array([0, 300, 300 + 601, 300 + 601 + 600, 300 + 601 + 600 + 2000], np.uint64)
    sr, ss, sl = oracle.segments(P, off)
    # 300 -> 1 segment; 601 > 2r -> 5 segments sampled over [0, len-r-1]; 600 -> 1 segment truncated to r; 2000 -> 5
    assert list(np.bincount(sr)) == [1, 5, 1, 5]
    assert list(ss[1:6]) == list(oracle.sample(5, 601 - 300 - 1))
    assert sl[6] == 300 and (sl[1:6] == 300).all()
    assert list(ss[7:12]) == list(oracle.sample(5, 2000 - 300 - 1))


def test_revcomp_roundtrip_through_map(oracle):
    """a read and its reverse complement give swapped candidate lists (q_gram_mapper.h:462-469) when every k-mer is
    sampled-position symmetric is not guaranteed, so only the weaker property is checked: forward candidates of a
    read from the forward strand contain its true bucket"""
    n_bases, L = 600_000, 65536
    packed = datagen.random_genome_packed(n_bases, 21)
    NB, bs, bl, br = datagen.bucketize(np.array([n_bases], np.uint64), L, 300)
    k2i, n_rows = oracle.fracminhash(9, 40503, 7919, 1.0)
    rows = oracle.index_build(packed, bs, bl, NB, 9, k2i, n_rows)
    seq, qual, off, start, rev = datagen.simulate_reads(packed, n_bases, 200, 300, 0.0, 0.0, 0.0, 22)
    P = oracle.params(L)
    segs = oracle.segments(P, off)
    coff, cand, stats = oracle.map_segments(P, NB, k2i, rows, seq, qual, off, segs)
    for i in range(200):
        lst = cand[coff[2 * i + int(rev[i])]:coff[2 * i + int(rev[i]) + 1]]
        assert int(start[i]) // L in lst or (int(start[i]) // L - 1 in lst and int(start[i]) % L < 300)


def test_zeros_and_phantom_buckets(oracle):
    # NB counts phantom buckets: an all-zero row has zeros == NB; threshold uses NB (q_gram_mapper.h:163,171-184)
    rows = np.zeros((3, 2), np.uint8)
    rows[1, 0] = 0b101
    rows[2, 0], rows[2, 1] = 0xFF, 0x1F            # all 13 buckets
    z, valid = oracle.zeros(rows, 13, 0.5)
    assert list(z) == [13, 11, 0]
    assert valid == 2                            # rows with zeros > (unsigned)(0.5 * 13) = 6
