"""GPU: the README accuracy table's metrics (mapped %, sensitivity, precision; sam_file_analyzer.cpp:250-358) for the
product command line on simulated reads with a ground truth file, through bucket-map_b200/bin/bucketmap_analyze.

A random multi-record genome has no repeats, so nearly every read must come back at its origin; a genome with duplicated
segments gives reads with two equally good locations, which the analyzer drops from its per-read counts (:306-314)."""
import json
import os
import subprocess

import numpy as np
import pytest

from bucketmap_b200 import datagen

pytestmark = [pytest.mark.gpu, pytest.mark.timeout(900)]

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, "bucket-map_b200", "bin")
L = 8192


def _map_and_analyze(d, packed, rec_len, n_reads, sub, indel, seed, align=False):
    n_bases = int(sum(rec_len))
    names = ["chr%d" % i for i in range(len(rec_len))]
    fa, fq, gt = os.path.join(d, "g.fa"), os.path.join(d, "r.fastq"), os.path.join(d, "r.position_ground_truth")
    datagen.write_fasta(fa, packed, rec_len, names)
    seq, qual, off, start, rev = datagen.simulate_reads(packed, n_bases, n_reads, 150, sub, indel, indel, seed)
    # reads that straddle two records have no single origin: keep those inside one record
    bounds = np.concatenate([[0], np.cumsum(rec_len)])
    rec = np.searchsorted(bounds, start, side="right") - 1
    inside = start + 160 <= bounds[rec + 1]
    keep = np.flatnonzero(inside)
    pieces = [(seq[off[i]:off[i + 1]], qual[off[i]:off[i + 1]]) for i in keep]
    koff = np.concatenate([[0], np.cumsum([len(p[0]) for p in pieces])]).astype(np.uint64)
    datagen.write_fastq(fq, np.concatenate([p[0] for p in pieces]), np.concatenate([p[1] for p in pieces]), koff)
    datagen.write_ground_truth(gt, start[keep], rev[keep], rec_len)
    exe = os.path.join(BIN, "bucketmap_align" if align else "bucketmap")
    for extra in (["-x"], ["-q", fq, "-o", "out.sam"]):
        p = subprocess.run([exe, "-i", "idx", "-g", fa, "-L", str(L), "-f", "1.0", "-r", "150"] + extra, cwd=d,
                           capture_output=True, text=True, timeout=600)
        assert p.returncode == 0, p.stderr[-2000:]
    p = subprocess.run([os.path.join(BIN, "bucketmap_analyze"), "--reads", fq, "--truth", gt, "--json",
                        os.path.join(d, "out.sam")], capture_output=True, text=True)
    assert p.returncode == 0, p.stderr
    return json.loads(p.stdout)[0], len(keep)


def test_random_genome_reads_come_back_at_their_origin(tmp_path):
    rec_len = [700_000, 180_000, 1_120_000]
    packed = datagen.random_genome_packed(sum(rec_len), 21)
    m, n = _map_and_analyze(str(tmp_path), packed, rec_len, 20_000, 0.002, 0.00025, 4)
    assert m["reads"] == n and m["answer_unique"] == n
    assert m["mapped_reads"] > 0.985 * n                # README: 98.6 % of the real reads mapped
    assert m["correct_reads"] > 0.98 * n                # README: 97.3 % at the right position
    assert m["acceptable_locations"] > 0.99 * m["mapped_locations"]


def test_align_stage_keeps_the_accuracy(tmp_path):
    rec_len = [900_000, 600_000]
    packed = datagen.random_genome_packed(sum(rec_len), 22)
    m, n = _map_and_analyze(str(tmp_path), packed, rec_len, 8_000, 0.01, 0.001, 5, align=True)
    assert m["mapped_reads"] > 0.9 * n
    assert m["correct_reads"] > 0.9 * n
    assert m["acceptable_locations"] > 0.99 * m["mapped_locations"]


def test_duplicated_segments_show_up_as_multiple_locations(tmp_path):
    n_bases = 2_000_000
    packed = datagen.add_repeats(datagen.random_genome_packed(n_bases, 23), n_bases, 5_000, 40, 6)
    m, n = _map_and_analyze(str(tmp_path), packed, [n_bases], 20_000, 0.0, 0.0, 7)
    multi = m["unique"][1] + m["unique"][3]             # reads the tested SAM has several records for
    assert multi > 0.03 * n                              # ~10 % of the genome is a copy of another part
    assert m["mapped_locations"] > m["mapped_reads"] + multi
    assert m["mapped_reads"] + multi > 0.985 * n
