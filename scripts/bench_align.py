#!/usr/bin/env python
"""C5 of BASELINE.json (bucketmap_align path): 1.7 Gbp synthetic reference, 1 M x 300 bp high-error reads (1 % sub,
0.1 % ins + 0.1 % del), map -> locate (no best-location merge) -> K4 bit-parallel verification with CIGAR.
Prints one JSON line: alignments/s and GCUPS of K4 (CUDA events around the kernels), reads/s of the whole path.
usage: python scripts/bench_align.py [--reads N] [--steps K]"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bucketmap_b200 as bm                      # noqa: E402
from bucketmap_b200 import datagen, hostlogic    # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--reads", type=int, default=1_000_000)
ap.add_argument("--steps", type=int, default=5)
a = ap.parse_args()
N_BASES, L, Q = 1_700_000_000, 65536, 9
ctx = bm.Context(0)
P = bm.default_params(ctx.lib, L)
packed = datagen.random_genome_packed(N_BASES, 3)
NB, b_start, b_len, _ = datagen.bucketize(np.array([N_BASES], np.uint64), L, 300)
k2i, n_rows = hostlogic.fracminhash(Q, 40503, 7919, 1.0)
ctx.ref_upload(packed, N_BASES, b_start, b_len, NB)
ctx.index_build(k2i, Q, n_rows, download=False)
seq, qual, off, start, rev = datagen.simulate_reads(packed, N_BASES, a.reads, 300, 0.01, 0.001, 0.001, 7)
ctx.reads_upload(P, seq, qual, off)
ctx.kernel_timing(True)


def step():
    ctx.map_device(P)
    ctx.locate_device(P, False)
    ctx.align_device(P)


step(); step()
ctx.sync()
ctx.kernel_stats_reset()
t0 = time.perf_counter()
ctx.timer_start()
for _ in range(a.steps):
    step()
ms = ctx.timer_stop()
wall = time.perf_counter() - t0
kst = ctx.kernel_stats()
loff, locs = ctx.locate_download()
aln, cig = ctx.align_download(len(locs))
n_aln = len(locs)
width = 300 + 1 + int(np.ceil(P.indel_rate * 300))     # text window columns (bucket_locator.h:548-552)
cells = n_aln * 300 * width
align_ms = kst["align"][0] / a.steps
first = loff[:-1]
has = np.diff(loff) > 0
best = np.full(len(first), -10 ** 9)
np.maximum.at(best, np.repeat(np.arange(len(first)), np.diff(loff)), aln["score"])
out = {
    "workload": "c5: 1.7 Gbp synthetic reference, L=65536, %d x 300bp reads (1%% sub, 0.1%% ins, 0.1%% del), bucketmap_align path" % a.reads,
    "reads_per_s": a.reads * a.steps / (ms * 1e-3), "ms_per_step": ms / a.steps, "wall_ms_per_step": wall * 1e3 / a.steps,
    "alignments_per_step": n_aln, "align_ms_per_step": align_ms,
    "alignments_per_s": n_aln / (align_ms * 1e-3), "gcups": cells / (align_ms * 1e-3) / 1e9,
    "cells_per_alignment": 300 * width,
    "kernels_ms_per_step": {k: v[0] / a.steps for k, v in kst.items() if v[1]},
    "reads_located": float(has.mean()), "mean_best_score": float(best[has].mean()),
    "kept": float(aln["kept"].mean()), "cigar_ops": int(len(cig)),
}
print(json.dumps(out))
ctx.close()
