"""dev: C2-scale mapper timing (1.7 Gbp, f=1).  python scripts/dev_c2.py [n_reads] [n_bases]"""
import sys, time, os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "oracle")); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import bucketmap_b200 as bm
from bucketmap_b200 import datagen
import bm_oracle

n_reads = int(sys.argv[1]) if len(sys.argv) > 1 else 200_000
n_bases = int(sys.argv[2]) if len(sys.argv) > 2 else 1_700_000_000
L = 65536
o = bm_oracle.Oracle()
t = time.time(); packed = datagen.random_genome_packed(n_bases, 3); print("genome", time.time() - t, flush=True)
NB, bs, bl, br = datagen.bucketize(np.array([n_bases], np.uint64), L, 300)
k2i, n_rows = o.fracminhash(9, 40503, 7919, 1.0)
ctx = bm.Context(0)
P = bm.default_params(ctx.lib, L)
ctx.kernel_timing(True)
t = time.time(); ctx.ref_upload(packed, n_bases, bs, bl, NB); print("ref upload", time.time() - t, flush=True)
t = time.time(); ctx.index_build(k2i, 9, n_rows, download=False); ctx.sync(); print("index build", time.time() - t, flush=True)
print(ctx.kernel_stats())
t = time.time(); seq, qual, off, ts, tr = datagen.simulate_reads(packed, n_bases, n_reads, 300, 0.002, 0.00025, 0.00025, 4); print("reads", time.time() - t, flush=True)
ctx.reads_upload(P, seq, qual, off)
for it in range(3):
    ctx.kernel_stats_reset()
    ctx.map_device(P); ctx.sync()
    st = ctx.kernel_stats()
    ms = st["map_vote"][0]
    goff, gcand, gstats = ctx.map_download()
    W = (NB + 7) >> 3
    bytes_alg = int(gstats[3]) * W
    print(f"iter {it}: map_vote {ms:.3f} ms  sample {st['map_sample'][0]:.3f} ms  rows {gstats[3]}  alg GB/s {bytes_alg / ms / 1e6:.1f}  reads/s {n_reads / ms * 1e3:.0f}  stats {gstats}", flush=True)
# parity on a prefix vs oracle
npre = 2000
rows = ctx.index_download()
Po = o.params(L)
sub_off = off[:npre + 1].copy()
segs = o.segments(Po, sub_off)
t = time.time(); coff, cand, stats = o.map_segments(Po, NB, k2i, rows, seq, qual, sub_off, segs); print("oracle map", time.time() - t, stats)
n2 = 2 * npre
assert np.array_equal(goff[:n2 + 1], coff), "cand_off mismatch"
assert np.array_equal(gcand[:coff[-1]], cand), "cand mismatch"
print("prefix parity OK")
