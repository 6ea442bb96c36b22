"""dev: C2 set-up once, then time the mapper / locator kernels for a sweep of K2f launch bounds ($BM_K2F_MINB) and with
the full-resolution table on / off.  usage: python scripts/dev_k2f.py [reads]"""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bucketmap_b200 as bm
from bucketmap_b200 import datagen, hostlogic

n_reads = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
N, L = 1_700_000_000, 65536
ctx = bm.Context(0)
P = bm.default_params(ctx.lib, L)
packed = datagen.random_genome_packed(N, 3)
NB, bs, bl, _ = datagen.bucketize(np.array([N], np.uint64), L, 300)
k2i, n_rows = hostlogic.fracminhash(9, 40503, 7919, 1.0)
ctx.ref_upload(packed, N, bs, bl, NB)
ctx.index_build(k2i, 9, n_rows, download=False)
seq, qual, off, _, _ = datagen.simulate_reads(packed, N, n_reads, 300, 0.002, 0.00025, 0.00025, 4)
ctx.reads_upload(P, seq, qual, off)
ctx.kernel_timing(True)
ref = None
for fine, minb, pfd, nod in ((1, "12", "3", "1"), (1, "12", "3", ""), (1, "14", "3", ""), (1, "12", "3", "1"), (1, "12", "3", "")):
    ctx.index_fine_table(fine)
    if True:
        os.environ["BM_K2F_MINB"] = minb
        os.environ["BM_K2F_LD"] = pfd
        if nod:
            os.environ["BM_K2F_NO_ALLDIST"] = "1"
        else:
            os.environ.pop("BM_K2F_NO_ALLDIST", None)
        for _ in range(2):
            ctx.map_device(P); ctx.locate_device(P, True)
        ctx.sync(); ctx.kernel_stats_reset()
        for _ in range(5):
            ctx.map_device(P); ctx.locate_device(P, True)
        ctx.sync()
        ks = ctx.kernel_stats()
        off_, locs = ctx.locate_download()
        if ref is None:
            ref = (off_.copy(), locs.copy())
        same = np.array_equal(ref[0], off_) and np.array_equal(ref[1], locs)
        print(f"fine={fine} minb={minb} ld={pfd} lookups={bool(nod)}: map {ks['map_vote'][0] / 5:.3f} ms  locate {ks['locate'][0] / 5:.3f}  aux {ks['aux'][0] / 5:.3f} "
              f"merge {ks['merge'][0] / 5:.3f}  same={same}", flush=True)
ctx.close()
