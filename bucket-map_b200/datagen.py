"""Seeded synthetic genomes and reads for tests and bench (the configs of BASELINE.json).

A deterministic restatement of what tools/short_read_simulator.h:40-240 does (the reference seeds it from
time()/random_device, so its output cannot be reproduced): a read is a window of a bucket, Poisson numbers of
deletions, insertions and substitutions are applied in that order (:114-116), the result is reverse-complemented
with probability 1/2 (:69-82) and written with all-'E' qualities (:225).  ACGT only.
"""
import numpy as np

_COMP = np.array([3, 2, 1, 0], np.uint8)
_ASCII = np.frombuffer(b"ACGT", np.uint8)


def random_genome_packed(n_bases, seed):
    """uniform i.i.d. genome, already 2-bit packed (base i in byte i>>2 at bits 2*(i&3)); + 16 zero pad bytes."""
    rng = np.random.default_rng(seed)
    nbytes = (n_bases + 3) // 4
    packed = np.frombuffer(rng.bytes(nbytes), np.uint8).copy()
    if n_bases & 3:
        packed[-1] &= (1 << (2 * (n_bases & 3))) - 1
    return np.concatenate([packed, np.zeros(16, np.uint8)])


def unpack(packed, start, n):
    """ranks (0..3) of bases [start, start+n)"""
    idx = np.arange(start, start + n, dtype=np.int64)
    return (packed[idx >> 2] >> ((idx & 3) << 1).astype(np.uint8)) & 3


def pack(ranks):
    n = len(ranks)
    pad = (-n) % 4
    r = np.concatenate([np.asarray(ranks, np.uint8), np.zeros(pad, np.uint8)]).reshape(-1, 4)
    packed = r[:, 0] | (r[:, 1] << 2) | (r[:, 2] << 4) | (r[:, 3] << 6)
    return np.concatenate([packed.astype(np.uint8), np.zeros(16, np.uint8)])


def add_repeats(packed, n_bases, seg_len, n_copies, seed):
    """copy random segments elsewhere so some reads have several candidate buckets (SURVEY 8d, C2')."""
    rng = np.random.default_rng(seed)
    ranks = unpack(packed, 0, n_bases).astype(np.uint8)
    for _ in range(n_copies):
        a = int(rng.integers(0, n_bases - seg_len))
        b = int(rng.integers(0, n_bases - seg_len))
        ranks[b:b + seg_len] = ranks[a:a + seg_len]
    return pack(ranks)


def bucketize(rec_len, L, r):
    """utils.h:60-102 -> (NB, b_start, b_len, b_rec).  NB follows CMakeLists.txt:13-50 (sum of ceil(len/L))."""
    b_start, b_len, b_rec = [], [], []
    base = 0
    NB = 0
    for rec, n in enumerate(rec_len):
        n = int(n)
        NB += (n + L - 1) // L
        nb = int(np.ceil(np.float32(n) / np.float32(L)))
        for i in range(nb):
            s = i * L
            e = min(s + L + r, n)
            if e - s <= r:
                continue
            b_start.append(base + s); b_len.append(e - s); b_rec.append(rec)
        base += n
    return NB, np.array(b_start, np.uint64), np.array(b_len, np.uint32), np.array(b_rec, np.uint32)


def simulate_reads(packed, n_bases, n_reads, read_len, sub_rate, ins_rate, del_rate, seed, chunk=65536, hot=None):
    """returns (seq ASCII uint8, qual ASCII uint8, read_off uint64[n+1], truth_start int64[n], truth_rev bool[n])"""
    rng = np.random.default_rng(seed)
    starts = rng.integers(0, n_bases - read_len - 1, n_reads, dtype=np.int64)
    if hot is not None:   # every 8th read starts inside [hot[0], hot[1])
        sel = np.arange(0, n_reads, 8)
        starts[sel] = rng.integers(max(0, hot[0]), min(hot[1], n_bases - read_len - 1), len(sel), dtype=np.int64)
    n_del = rng.poisson(del_rate * read_len, n_reads)
    n_ins = rng.poisson(ins_rate * read_len, n_reads)
    n_sub = rng.poisson(sub_rate * read_len, n_reads)
    rev = rng.integers(0, 2, n_reads).astype(bool)
    lens = read_len - n_del + n_ins
    lens = np.maximum(lens, 1)
    off = np.zeros(n_reads + 1, np.uint64)
    np.cumsum(lens, out=off[1:])
    seq = np.empty(int(off[-1]), np.uint8)
    ar = np.arange(read_len, dtype=np.int64)
    for c0 in range(0, n_reads, chunk):
        c1 = min(c0 + chunk, n_reads)
        idx = starts[c0:c1, None] + ar[None, :]
        win = ((packed[idx >> 2] >> ((idx & 3) << 1).astype(np.uint8)) & 3).astype(np.uint8)
        simple = (n_del[c0:c1] == 0) & (n_ins[c0:c1] == 0)
        # substitutions on the simple reads, vectorised: pick positions, add 1..3 mod 4
        for t in range(int(n_sub[c0:c1].max(initial=0))):
            rows = np.nonzero(simple & (n_sub[c0:c1] > t))[0]
            if len(rows) == 0:
                continue
            pos = rng.integers(0, read_len, len(rows))
            win[rows, pos] = (win[rows, pos] + rng.integers(1, 4, len(rows)).astype(np.uint8)) & 3
        rv = rev[c0:c1]
        out = win.copy()
        out[rv] = _COMP[win[rv][:, ::-1]]
        o = off[c0:c1 + 1].astype(np.int64)
        if simple.all():
            seq[o[0]:o[-1]] = out.reshape(-1)
        else:
            for i in range(c1 - c0):
                if simple[i]:
                    seq[o[i]:o[i + 1]] = out[i]
                    continue
                s = list(win[i])
                for _ in range(int(n_del[c0 + i])):
                    if len(s) > 1:
                        del s[int(rng.integers(0, len(s)))]
                for _ in range(int(n_ins[c0 + i])):
                    s.insert(int(rng.integers(0, len(s))), int(rng.integers(0, 4)))
                for _ in range(int(n_sub[c0 + i])):
                    p = int(rng.integers(0, len(s)))
                    s[p] = (s[p] + int(rng.integers(1, 4))) & 3
                a = np.array(s, np.uint8)
                if rev[c0 + i]:
                    a = _COMP[a[::-1]]
                assert len(a) == o[i + 1] - o[i]
                seq[o[i]:o[i + 1]] = a
    seq = _ASCII[seq]
    qual = np.full(len(seq), ord("E"), np.uint8)
    return seq, qual, off, starts, rev


def write_fasta(path, packed, rec_len, names, width=60):
    with open(path, "wb") as f:
        base = 0
        for n, name in zip(rec_len, names):
            f.write(b">" + name.encode() + b"\n")
            s = _ASCII[unpack(packed, base, int(n))]
            for i in range(0, len(s), width):
                f.write(s[i:i + width].tobytes() + b"\n")
            base += int(n)


def write_fastq(path, seq, qual, off, names=None):
    with open(path, "wb") as f:
        for i in range(len(off) - 1):
            a, b = int(off[i]), int(off[i + 1])
            nm = str(i) if names is None else names[i]
            f.write(b"@" + nm.encode() + b"\n" + seq[a:b].tobytes() + b"\n+\n" + qual[a:b].tobytes() + b"\n")


def write_fasta_fast(path, packed, n_bases, name, width=60, chunk=1 << 26):
    """single-record FASTA of a packed genome, vectorised (the per-line loop of write_fasta is too slow at Gbp scale)"""
    lut = np.zeros((256, 4), np.uint8)
    v = np.arange(256)
    for j in range(4):
        lut[:, j] = _ASCII[(v >> (2 * j)) & 3]
    chunk -= chunk % (4 * width)
    with open(path, "wb") as f:
        f.write(b">" + name.encode() + b"\n")
        for b0 in range(0, n_bases, chunk):
            n = min(chunk, n_bases - b0)
            s = lut[packed[b0 // 4:(b0 + n + 3) // 4]].reshape(-1)[:n]
            full = n - n % width
            if full:
                lines = np.empty((full // width, width + 1), np.uint8)
                lines[:, :width] = s[:full].reshape(-1, width)
                lines[:, width] = 10
                f.write(lines.tobytes())
            if n % width:
                f.write(s[full:].tobytes() + b"\n")


def write_fastq_fast(path, seq, qual, off):
    """FASTQ with numeric read names; vectorised for equal-length reads, per-read loop otherwise"""
    n = len(off) - 1
    lens = np.diff(off.astype(np.int64))
    if n and (lens == lens[0]).all():
        m = int(lens[0])
        names = np.char.add(np.char.add("@", np.arange(n).astype(str)), "\n").astype("S")
        body = np.empty((n, 2 * m + 4), np.uint8)
        body[:, :m] = seq.reshape(n, m)
        body[:, m] = 10; body[:, m + 1] = ord("+"); body[:, m + 2] = 10
        body[:, m + 3:2 * m + 3] = qual.reshape(n, m)
        body[:, 2 * m + 3] = 10
        with open(path, "wb") as f:
            for i in range(n):
                f.write(names[i]); f.write(body[i].tobytes())
        return
    write_fastq(path, seq, qual, off)


def add_repeats_packed(packed, n_bases, seg_len, n_copies, seed):
    """add_repeats in the packed domain (4-base aligned source / destination / length), for Gbp genomes: copies
    n_copies random segments of seg_len bases elsewhere (SURVEY 8d, C2': 5 % duplicated 10 kbp segments)"""
    rng = np.random.default_rng(seed)
    out = packed.copy()
    nb = seg_len // 4
    lim = (n_bases - seg_len) // 4
    src = rng.integers(0, lim, n_copies)
    dst = rng.integers(0, lim, n_copies)
    for a, b in zip(src.tolist(), dst.tolist()):
        out[b:b + nb] = out[a:a + nb]
    return out


def add_family_packed(packed, n_bases, unit_len, n_copies, divergence, seed):
    """a high-copy repeat family (Alu-like): n_copies copies of ONE random element of unit_len bases, each with its own
    substitutions (rate `divergence`), written at random 4-base-aligned positions.  Reads that fall inside a copy share
    their k-mers with copies in almost every bucket: the mapper's worst case (coarse vote passes everywhere, the list goes
    to the full-width gather and comes back longer than 30 buckets, i.e. empty, in the reference as here)."""
    rng = np.random.default_rng(seed)
    out = packed.copy()
    unit_len -= unit_len % 4
    unit = rng.integers(0, 4, unit_len).astype(np.uint8)
    pos = rng.integers(0, (n_bases - unit_len) // 4, n_copies)
    chunk = 4096
    for c0 in range(0, n_copies, chunk):
        c1 = min(c0 + chunk, n_copies)
        copies = np.broadcast_to(unit, (c1 - c0, unit_len)).copy()
        mut = rng.random(copies.shape) < divergence
        copies[mut] = (copies[mut] + rng.integers(1, 4, int(mut.sum())).astype(np.uint8)) & 3
        r = copies.reshape(c1 - c0, unit_len // 4, 4)
        by = (r[:, :, 0] | (r[:, :, 1] << 2) | (r[:, :, 2] << 4) | (r[:, :, 3] << 6)).astype(np.uint8)
        for i, p in enumerate(pos[c0:c1].tolist()):
            out[p:p + unit_len // 4] = by[i]
    return out


def write_ground_truth(path, starts, rev, rec_len=None):
    """the simulator's `.position_ground_truth` (tools/short_read_simulator.h:237-240): one line per read,
    `<record index> <1-based position inside the record> <reverse complement 0|1> <cigar>`; this generator keeps no
    CIGAR of its edits, the column (never read by the analyzer, sam_file_analyzer.cpp:141) is written as `*`."""
    starts = np.asarray(starts, np.int64)
    if rec_len is None:
        rec = np.zeros(len(starts), np.int64)
        pos = starts
    else:
        bounds = np.concatenate([[0], np.cumsum(np.asarray(rec_len, np.int64))])
        rec = np.searchsorted(bounds, starts, side="right") - 1
        pos = starts - bounds[rec]
    with open(path, "w") as f:
        f.write("".join("%d %d %d *\n" % (r, p + 1, v) for r, p, v in zip(rec.tolist(), pos.tolist(),
                                                                         np.asarray(rev, bool).astype(int).tolist())))
