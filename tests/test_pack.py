"""CPU: the host-side read packer of the C-ABI (bm_pack_layout / bm_pack_reads, the FASTQ parser's output stage)
against a numpy restatement: 2-bit base planes, 32 bases per chunk, and the k-mer quality plane
(views::kmer_quality window sums >= b*k, mapper/quality_filter.h:531-534)."""
import numpy as np
import pytest


def np_pack(seq, qual, off, k, b):
    rank = np.zeros(256, np.uint8)
    for ch, r in (("C", 1), ("Y", 1), ("S", 1), ("B", 1), ("G", 2), ("K", 2), ("T", 3), ("U", 3)):
        rank[ord(ch)] = r
        rank[ord(ch.lower())] = r
    n = len(off) - 1
    lens = np.diff(off.astype(np.int64))
    nch = (lens + 31) // 32
    pl_off = np.zeros(n + 1, np.uint32)
    pl_off[1:] = np.cumsum(nch)
    planes = np.zeros(int(pl_off[-1]) + 1, np.uint64)
    qk = np.zeros(int(pl_off[-1]) + 1, np.uint32)
    low = 0
    for i in range(n):
        s = rank[seq[off[i]:off[i + 1]]].astype(np.uint64)
        q = qual[off[i]:off[i + 1]].astype(np.int64) - 33
        m = len(s)
        for c in range(int(nch[i])):
            seg = s[32 * c:32 * c + 32]
            t = np.arange(len(seg), dtype=np.uint64)
            lo = int(((seg & 1) << t).sum())
            hi = int(((seg >> 1) << t).sum())
            planes[pl_off[i] + c] = lo | (hi << 32)
        nk = max(0, m - k + 1)
        fail = False
        for j in range(nk):
            ok = q[j:j + k].sum() >= b * k
            fail |= not ok
            if ok:
                qk[pl_off[i] + (j >> 5)] |= np.uint32(1 << (j & 31))
        low += int(fail)
    return planes, qk, pl_off, low


@pytest.mark.parametrize("k,b", [(12, 25), (9, 30), (16, 2), (32, 25)])
def test_pack_matches_numpy(bm, k, b):
    lib = bm.load_library()
    rng = np.random.default_rng(k * 100 + b)
    lens = np.concatenate([rng.integers(1, 400, 60), [0, 1, 31, 32, 33, 63, 64, 65, k - 1, k, k + 1, 300, 300]])
    off = np.zeros(len(lens) + 1, np.uint64)
    off[1:] = np.cumsum(lens)
    alphabet = np.frombuffer(b"ACGTacgtNnRYKMSWBDHVU", np.uint8)
    seq = alphabet[rng.integers(0, len(alphabet), int(off[-1]))]
    seq[: int(off[30])] = alphabet[rng.integers(0, 4, int(off[30]))]
    qual = np.full(int(off[-1]), ord("E"), np.uint8)
    lowq = rng.random(int(off[-1])) < 0.08
    qual[lowq] = rng.integers(33, 33 + 30, int(lowq.sum()))
    qual[: int(off[20])] = ord("E")                       # the first reads are clean: all windows pass
    planes, qk, pl_off, low = bm.pack_reads(lib, seq, qual, off, k, b)
    eplanes, eqk, epl_off, elow = np_pack(seq, qual, off, k, b)
    assert np.array_equal(pl_off, epl_off)
    assert np.array_equal(planes, eplanes)
    assert np.array_equal(qk, eqk)
    assert low == elow


def test_pack_ranges_compose(bm):
    """several threads pack disjoint read ranges into the same arrays"""
    import ctypes as C
    lib = bm.load_library()
    rng = np.random.default_rng(5)
    lens = rng.integers(20, 310, 200)
    off = np.zeros(len(lens) + 1, np.uint64)
    off[1:] = np.cumsum(lens)
    seq = np.frombuffer(b"ACGT", np.uint8)[rng.integers(0, 4, int(off[-1]))]
    qual = rng.integers(33 + 10, 33 + 41, int(off[-1])).astype(np.uint8)
    planes, qk, pl_off, low = bm.pack_reads(lib, seq, qual, off, 12, 25)
    p2 = np.zeros_like(planes)
    q2 = np.zeros_like(qk)
    tot = 0
    for r0, r1 in ((0, 50), (50, 51), (51, 200)):
        nl = C.c_uint32()
        rc = lib.bm_pack_reads(seq.ctypes.data, qual.ctypes.data, off.ctypes.data, pl_off.ctypes.data, r0, r1, 12, 25,
                               p2.ctypes.data, q2.ctypes.data, C.byref(nl))
        assert rc == 0
        tot += nl.value
    assert np.array_equal(p2, planes) and np.array_equal(q2, qk) and tot == low
