#!/usr/bin/env python
"""Wall time of the drop-in command line on C2: FASTA (1.7 Gbp) + FASTQ (1 M x 300 bp) in, SAM out, through
bucket-map_b200/bin/bucketmap exactly as a user of the reference would run it (`-x` then `-q/-o`).
Prints one JSON line.  usage: python scripts/bench_cli.py [--reads N] [--dir D]"""
import argparse
import json
import os
import re
import subprocess
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from bucketmap_b200 import datagen   # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--reads", type=int, default=1_000_000)
ap.add_argument("--dir", default=None)
ap.add_argument("--bases", type=int, default=1_700_000_000)
a = ap.parse_args()
d = a.dir or tempfile.mkdtemp(prefix="bmcli_", dir="/dev/shm" if os.path.isdir("/dev/shm") else None)
exe = os.path.join(ROOT, "bucket-map_b200", "bin", "bucketmap")
packed = datagen.random_genome_packed(a.bases, 3)
seq, qual, off, start, rev = datagen.simulate_reads(packed, a.bases, a.reads, 300, 0.002, 0.00025, 0.00025, 4)
fa, fq = os.path.join(d, "genome.fa"), os.path.join(d, "reads.fq")
datagen.write_fasta_fast(fa, packed, a.bases, "chr1")
datagen.write_fastq_fast(fq, seq, qual, off)


def run(args):
    t = time.perf_counter()
    p = subprocess.run([exe] + args, cwd=d, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True)
    dt = time.perf_counter() - t
    if p.returncode:
        sys.stderr.write(p.stderr[-4000:])
        sys.exit(1)
    return dt, p.stderr


common = ["-g", fa, "-L", "65536", "-f", "1", "--seed-x", "40503", "--seed-y", "7919"]
t_index, log_x = run(["-x", "-i", "idx"] + common)
t_map, log_m = run(["-i", "idx", "-q", fq, "-o", "out.sam"] + common)
sam_bytes = os.path.getsize(os.path.join(d, "out.sam"))
bench_lines = [l for l in log_m.splitlines() if l.startswith("[BENCHMARK]") or l.startswith("[pipe]") or l.startswith("[host]")]
m = re.search(r"Total mapped locations: (\d+)", log_m)
out = {"workload": "c2 through the command line: genome.fa %d bases, reads.fq %d x 300bp" % (a.bases, a.reads),
       "dir": d, "index_wall_s": t_index, "map_wall_s": t_map, "reads_per_s_map_wall": a.reads / t_map,
       "fastq_bytes": os.path.getsize(fq), "sam_bytes": sam_bytes, "locations": int(m.group(1)) if m else None,
       "tool_log": bench_lines}
print(json.dumps(out))
for f in ("genome.fa", "reads.fq", "out.sam", "idx.qgram", "idx.kmers_index", "idx.bucket_id"):
    try:
        os.remove(os.path.join(d, f))
    except OSError:
        pass
