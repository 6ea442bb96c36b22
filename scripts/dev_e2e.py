"""dev: phases of bm_map_locate_packed on C2 ($BM_PIPE_TRACE=1 prints them per call).  usage: python scripts/dev_e2e.py [--plain]"""
import os
import sys
import time

import numpy as np

if "--plain" not in sys.argv:
    os.environ["BM_PIPE_TRACE"] = "1"
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bucketmap_b200 as bm
from bucketmap_b200 import datagen, hostlogic

N, L = 1_700_000_000, 65536
n_reads = 1_000_000
ctx = bm.Context(0)
P = bm.default_params(ctx.lib, L)
packed = datagen.random_genome_packed(N, 3)
NB, bs, bl, _ = datagen.bucketize(np.array([N], np.uint64), L, 300)
k2i, n_rows = hostlogic.fracminhash(9, 40503, 7919, 1.0)
ctx.ref_upload(packed, N, bs, bl, NB)
ctx.index_build(k2i, 9, n_rows, download=False)
seq, qual, off, _, _ = datagen.simulate_reads(packed, N, n_reads, 300, 0.002, 0.00025, 0.00025, 4)
planes, qk, pl_off, low = bm.pack_reads(ctx.lib, seq, qual, off, P.k, P.avg_base_quality)
hpl = ctx.pinned(len(planes), np.uint64); hpl[:] = planes
hplo = ctx.pinned(len(pl_off), np.uint32); hplo[:] = pl_off
hoff = ctx.pinned(len(off), np.uint64); hoff[:] = off
ooff = ctx.pinned(n_reads + 1, np.uint32)
olocs = ctx.pinned(4 * n_reads, bm.LOC_DTYPE)
for i in range(10 if "--plain" in sys.argv else 6):
    t = time.perf_counter()
    ctx.map_locate_packed_into(P, hpl, None, hplo, hoff, ooff, olocs, True)
    print("call %d: %.2f ms" % (i, (time.perf_counter() - t) * 1e3), flush=True)
if "--plain" in sys.argv:
    # two batches in flight: this context and a lane, one host thread each
    import threading
    lane = ctx.create_lane(P)
    ooff2 = ctx.pinned(n_reads + 1, np.uint32)
    olocs2 = ctx.pinned(4 * n_reads, bm.LOC_DTYPE)
    steps = 10

    def drive(c, oo, ol):
        for _ in range(steps):
            c.map_locate_packed_into(P, hpl, None, hplo, hoff, oo, ol, True)

    for rep in range(3):
        th = [threading.Thread(target=drive, args=a) for a in ((ctx, ooff, olocs), (lane, ooff2, olocs2))]
        t = time.perf_counter()
        for x in th:
            x.start()
        for x in th:
            x.join()
        dt = time.perf_counter() - t
        print("two lanes: %.2f ms per batch  (%.1f M reads/s)  same=%s" % (
            dt / (2 * steps) * 1e3, 2 * steps * n_reads / dt / 1e6,
            np.array_equal(ooff, ooff2) and np.array_equal(olocs[:int(ooff[-1])], olocs2[:int(ooff2[-1])])), flush=True)
    lane.close()
ctx.close()
