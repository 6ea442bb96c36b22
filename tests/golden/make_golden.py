#!/usr/bin/env python
"""Generates tests/golden/*: outputs of the REFERENCE ITSELF on small seeded inputs.

The reference ships no golden vectors, so the oracle is pinned against the reference's own code
run here: /root/reference/bucket_map/main.cpp compiled unmodified against oracle/shim (oracle/Makefile `ref`), executed
as `bucketmap -x -i` then `bucketmap[_align] -i -q -o` on each case below.  Needs /root/reference (this container);
the committed files do not.  Per case:

  <case>.npz        inputs: 2-bit packed genome, record lengths/names, reads (ASCII bases, qualities, offsets, names)
  <case>.json       parameters, NB/L, FracMinHash (x, y), sha256 of <ind>.qgram / <ind>.kmers_index, the .bucket_id
                    text, the reference's own log counters
  <case>.sam.gz     SAM written by the reference's bucketmap
  <case>.align.sam.gz   SAM written by the reference's bucketmap_align (cases with align=True)

Run:  python tests/golden/make_golden.py [case ...]
"""
import gzip
import hashlib
import json
import os
import shutil
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))

import bm_ref                                   # noqa: E402
from bucketmap_b200 import datagen              # noqa: E402  (seeded data generator only; no kernels)

X, Y = 40503, 7919

# name -> dict(genome=..., reads=..., flags=..., align=bool)
CASES = {
    # defaults (-f 0.25), 150 bp reads, one record
    "short150": dict(n_bases=400_000, L=8192, rec_split=1, n_reads=400, read_len=150, sub=0.002, indel=0.00025,
                     flags=dict(), align=True),
    # -f 1 (every q-gram indexed: the roofline configuration), 300 bp reads
    "full300": dict(n_bases=400_000, L=8192, rec_split=1, n_reads=300, read_len=300, sub=0.002, indel=0.00025,
                    flags=dict(f=1.0), align=True),
    # several FASTA records with descriptions; one record shorter than -r (skipped bucket => phantom bucket ids)
    "multirec": dict(n_bases=500_000, L=8192, rec_split=6, short_rec=200, n_reads=300, read_len=300, sub=0.002,
                     indel=0.00025, flags=dict(f=1.0), align=False),
    # mixed / low base qualities: quality filter, all-k-mers fallback of the locator, too-few-good-k-mers bail-out
    "lowqual": dict(n_bases=300_000, L=8192, rec_split=1, n_reads=300, read_len=300, sub=0.002, indel=0.00025,
                    flags=dict(f=1.0), lowqual=True, align=False),
    # reads longer than 2r: five sampled segments merged by _filter_best_locations
    "long800": dict(n_bases=400_000, L=8192, rec_split=1, n_reads=120, read_len=800, sub=0.002, indel=0.00025,
                    flags=dict(f=1.0), align=False),
    # duplicated segments: several candidate buckets per read, vote merges
    "repeats": dict(n_bases=400_000, L=8192, rec_split=1, repeats=30, n_reads=300, read_len=300, sub=0.002,
                    indel=0.00025, flags=dict(f=1.0), align=True),
    # high-error reads with the benchmark flags of benchmark/short_read/benchmark_map.sh (-s 20 -e 0.6 -l 14 -b 10)
    "higherr": dict(n_bases=400_000, L=8192, rec_split=1, n_reads=300, read_len=300, sub=0.01, indel=0.001,
                    flags=dict(f=1.0, s=20, e=0.6, l=14, b=10), align=True),
    # larger bucket length, ragged read lengths (100..300)
    "ragged": dict(n_bases=2_400_000, L=65536, rec_split=2, n_reads=300, read_len=300, sub=0.004, indel=0.002,
                   flags=dict(), ragged=True, align=True),
}


def make_inputs(c, seed):
    n = c["n_bases"]
    packed = datagen.random_genome_packed(n, seed)
    if c.get("repeats"):
        packed = datagen.add_repeats(packed, n, 3000, c["repeats"], seed + 100)
    if c["rec_split"] == 1:
        rec_len = np.array([n], np.uint64)
    else:
        rng = np.random.default_rng(seed + 7)
        cuts = np.sort(rng.integers(1000, n - 1000, c["rec_split"] - 1))
        if c.get("short_rec"):
            cuts[1] = cuts[0] + c["short_rec"]          # record 1 is shorter than -r: its only bucket is skipped
            cuts = np.sort(cuts)
        rec_len = np.diff(np.concatenate([[0], cuts, [n]])).astype(np.uint64)
    names = [f"chr{i + 1} synthetic record {i + 1} len={int(l)}" for i, l in enumerate(rec_len)]
    seq, qual, off, _, _ = datagen.simulate_reads(packed, n, c["n_reads"], c["read_len"], c["sub"], c["indel"],
                                                 c["indel"], seed + 1)
    rng = np.random.default_rng(seed + 9)
    if c.get("ragged"):                                   # cut every read to a random length in [100, len]
        keep_seq, keep_off = [], [0]
        for i in range(len(off) - 1):
            a, b = int(off[i]), int(off[i + 1])
            m = int(rng.integers(100, b - a + 1))
            keep_seq.append(seq[a:a + m]); keep_off.append(keep_off[-1] + m)
        seq = np.concatenate(keep_seq); off = np.array(keep_off, np.uint64)
        qual = np.full(len(seq), ord("E"), np.uint8)
    if c.get("lowqual"):
        qual = qual.copy()
        for i in range(len(off) - 1):
            a, b = int(off[i]), int(off[i + 1])
            if i % 3 == 1:
                qual[a:b] = (33 + rng.integers(0, 42, b - a)).astype(np.uint8)       # mixed
            elif i % 9 == 2:
                qual[a:b] = 33 + 5                                                   # all low
            elif i % 9 == 5:
                qual[a:b] = 33 + 5; qual[a + 40:a + 40 + 14] = 33 + 40                # a single good window
    read_names = [f"read{i} case" if i % 5 == 0 else f"read{i}" for i in range(len(off) - 1)]
    return packed, rec_len, names, seq, qual, off, read_names


def sha(path):
    return hashlib.sha256(open(path, "rb").read()).hexdigest()


def run_case(name, c, seed):
    packed, rec_len, names, seq, qual, off, read_names = make_inputs(c, seed)
    L = c["L"]
    NB = int(sum((int(n) + L - 1) // L for n in rec_len))
    bm_ref.build(NB, L)
    work = tempfile.mkdtemp(prefix="bm_golden_")
    try:
        fa, fq = os.path.join(work, "genome.fa"), os.path.join(work, "reads.fq")
        datagen.write_fasta(fa, packed, rec_len, names)
        datagen.write_fastq(fq, seq, qual, off, read_names)
        flags = dict(c["flags"])
        bm_ref.index(NB, L, work, fa, "idx", X, Y, **flags)
        meta = dict(case=name, seed=seed, NB=NB, L=L, x=X, y=Y, flags=flags, n_bases=int(c["n_bases"]),
                    n_reads=len(off) - 1,
                    qgram_sha256=sha(os.path.join(work, "idx.qgram")),
                    qgram_bytes=os.path.getsize(os.path.join(work, "idx.qgram")),
                    kmers_index_sha256=sha(os.path.join(work, "idx.kmers_index")),
                    bucket_id=open(os.path.join(work, "idx.bucket_id")).read())
        sam = os.path.join(work, "out.sam")
        log = bm_ref.map_reads(NB, L, work, fa, "idx", fq, sam, X, Y, align=False, **flags)
        meta["ref_counters"] = {k: v for k, v in log.items() if k in ("with_candidate", "locations")}
        with gzip.GzipFile(os.path.join(HERE, name + ".sam.gz"), "wb", mtime=0) as f:
            f.write(open(sam, "rb").read())
        if c["align"]:
            sam2 = os.path.join(work, "out_align.sam")
            log2 = bm_ref.map_reads(NB, L, work, fa, "idx", fq, sam2, X, Y, align=True, **flags)
            meta["ref_counters_align"] = {k: v for k, v in log2.items() if k in ("with_candidate", "locations")}
            with gzip.GzipFile(os.path.join(HERE, name + ".align.sam.gz"), "wb", mtime=0) as f:
                f.write(open(sam2, "rb").read())
        np.savez_compressed(os.path.join(HERE, name + ".npz"), packed=packed, rec_len=rec_len,
                            names=np.array(names), seq=seq, qual=qual, read_off=off, read_names=np.array(read_names))
        json.dump(meta, open(os.path.join(HERE, name + ".json"), "w"), indent=1, sort_keys=True)
        print(f"{name}: NB={NB} reads={len(off) - 1} {meta['ref_counters']}"
              + (f" align {meta['ref_counters_align']}" if c["align"] else ""), flush=True)
    finally:
        shutil.rmtree(work, ignore_errors=True)


if __name__ == "__main__":
    todo = sys.argv[1:] or list(CASES)
    for i, (name, c) in enumerate(CASES.items()):
        if name in todo:
            run_case(name, c, 100 + 10 * i)
