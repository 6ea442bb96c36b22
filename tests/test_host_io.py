"""CPU: the host I/O stage of the C++ side (bucket-map_b200/host/bm_io.cpp) -- the multi-threaded FASTA packer (utils.h:60-102:
records, 2-bit text, bucket table) and the FASTQ splitter / packer -- against numpy, for 1 and several host threads."""
import os
import subprocess

import numpy as np
import pytest

import bucketmap_b200 as bm
from bucketmap_b200 import datagen

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "bucket-map_b200")


def fnv(data):
    h = 1469598103934665603
    for b in bytes(data):
        h = ((h ^ b) * 1099511628211) & 0xFFFFFFFFFFFFFFFF
    return "%016x" % h


def fnv_np(a):
    return fnv(np.ascontiguousarray(a).tobytes())


@pytest.fixture(scope="module")
def harness(tmp_path_factory):
    out = str(tmp_path_factory.mktemp("ioh") / "io_harness")
    host = os.path.join(PKG, "host")
    subprocess.check_call(["g++", "-O2", "-std=c++17", "-I", os.path.join(ROOT, "include"), "-I", host, "-o", out,
                           os.path.join(ROOT, "tests", "host", "io_harness.cpp"), os.path.join(host, "bm_io.cpp"),
                           "-L", PKG, "-lbucketmap_b200", "-Wl,-rpath," + PKG, "-pthread"])
    return out


def _fasta_text(rng, recs, width, crlf, junk_before):
    """recs: list of (name, ranks).  Returns bytes with the chosen line ending; some lines carry blanks / tabs / lower case."""
    nl = b"\r\n" if crlf else b"\n"
    out = []
    if junk_before:
        out.append(b"ACGTACGTTT" + nl)                         # bases before the first header belong to no record
    for name, ranks in recs:
        out.append(b">" + name.encode() + nl)
        s = np.frombuffer(b"ACGT", np.uint8)[ranks].copy()
        low = rng.random(len(s)) < 0.1
        s[low] |= 0x20                                         # lower case
        n_mask = rng.random(len(s)) < 0.01
        s[n_mask] = ord("N")                                   # N -> A (dna5 -> dna4)
        for i in range(0, len(s), width):
            line = s[i:i + width].tobytes()
            if rng.random() < 0.05 and len(line) > 10:
                line = line[:5] + b" \t" + line[5:]
            out.append(line + nl)
        ranks[n_mask] = 0
    text = b"".join(out)
    return text[:-len(nl)] if rng.random() < 0.5 else text     # with and without a final newline


@pytest.mark.parametrize("crlf,junk,threads", [(False, False, 1), (False, True, 5), (True, False, 7), (True, True, 3)])
def test_fasta_packer_matches_numpy(harness, tmp_path, crlf, junk, threads):
    rng = np.random.default_rng(17 + threads)
    # big enough for the multi-threaded path (> 4 MiB of text), several records incl. an empty and a tiny one
    lens = [1_700_003, 0, 250, 2_900_001, 77_777]
    recs = [("chr%d some description %d" % (i, i), rng.integers(0, 4, n).astype(np.uint8)) for i, n in enumerate(lens)]
    fa = tmp_path / "g.fa"
    fa.write_bytes(_fasta_text(rng, recs, 61, crlf, junk))
    L, r = 65536, 300
    p = subprocess.run([harness, "fasta", str(fa), str(L), str(r), str(threads)], stdout=subprocess.PIPE, text=True, check=True)
    lines = p.stdout.splitlines()
    ranks = np.concatenate([x[1] for x in recs])
    NB, b_start, b_len, b_rec = datagen.bucketize(np.array(lens, np.uint64), L, r)
    assert lines[0] == "n_bases %d NB %d records %d buckets %d" % (len(ranks), NB, len(lens), len(b_start))
    for i, (name, _) in enumerate(recs):
        assert lines[1 + i] == "rec %d %s" % (lens[i], name)
    packed = datagen.pack(ranks)[:(len(ranks) + 3) // 4]
    assert lines[1 + len(recs)] == "packed " + fnv_np(packed)
    got = [tuple(int(x) for x in ln.split()[1:]) for ln in lines[2 + len(recs):]]
    assert got == list(zip(b_start.tolist(), b_len.tolist(), b_rec.tolist()))


@pytest.mark.parametrize("threads,batch,crlf", [(1, 100000, False), (6, 100000, False), (4, 777, False), (5, 1500, True)])
def test_fastq_splitter_and_packer_match_numpy(harness, tmp_path, threads, batch, crlf):
    nl = b"\r\n" if crlf else b"\n"
    lib = bm.load_library()
    rng = np.random.default_rng(5)
    n = 6000
    lens = rng.integers(30, 320, n)
    off = np.zeros(n + 1, np.uint64)
    off[1:] = np.cumsum(lens)
    seq = np.frombuffer(b"ACGTacgtN", np.uint8)[rng.integers(0, 9, int(off[-1]))]
    qual = rng.integers(33 + 2, 33 + 41, int(off[-1])).astype(np.uint8)
    qual[: int(off[3000])] = ord("I")                            # the first half is clean
    short = set(rng.choice(n, 40, replace=False).tolist())       # quality line shorter than the read: padded with '!'
    lines = []
    ids = []
    qual_eff = qual.copy()
    for i in range(n):
        a, b = int(off[i]), int(off[i + 1])
        name = b"read_%d extra words" % i
        ids.append(name)
        q = qual[a:b].tobytes()
        if i in short:
            q = q[:len(q) // 2]
            qual_eff[a + len(q):b] = ord("!")
        lines.append(b"@" + name + nl + seq[a:b].tobytes() + nl + b"+" + nl + q + nl)
        if i % 1000 == 7:
            lines.append(nl)                                  # stray empty line between records
    fq = tmp_path / "r.fq"
    fq.write_bytes(b"".join(lines))
    k, bq = 12, 25
    p = subprocess.run([harness, "fastq", str(fq), str(batch), str(k), str(bq), str(threads)], stdout=subprocess.PIPE, text=True,
                       check=True)
    out = p.stdout.splitlines()
    assert len(out) == (n + batch - 1) // batch
    r0 = 0
    for ln in out:
        f = ln.split()
        m = int(f[1])
        lo, hi = r0, r0 + m
        sub_off = (off[lo:hi + 1] - off[lo]).astype(np.uint64)
        a, b = int(off[lo]), int(off[hi])
        planes, qk, pl_off, low = bm.pack_reads(lib, seq[a:b], qual_eff[a:b], sub_off, k, bq)
        assert int(f[3]) == int(pl_off[-1])
        assert int(f[5]) == (0 if low else 1)
        assert f[7] == fnv_np(planes[:pl_off[-1]])
        assert f[9] == (fnv_np(qk[:pl_off[-1]]) if low else "%016x" % 0)    # the plane exists only for a batch that needs it
        assert f[11] == fnv_np(sub_off) and f[13] == fnv_np(pl_off)
        assert f[15] == fnv(b"".join(ids[lo:hi]))
        r0 = hi
    assert r0 == n
