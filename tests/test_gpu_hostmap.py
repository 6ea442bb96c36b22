"""GPU: the mapper interface of the C++ boundary on its own -- bm::q_gram_mapper::load / map() / reset
(mapper/mapper.h:12-33, q_gram_mapper.h:318-378,483-557): bucket-major inverted lists of (read id, segment start) in read
order, and num_records, against the CPU oracle."""
import os
import subprocess

import numpy as np
import pytest

from golden_util import Golden

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "bucket-map_b200")


@pytest.fixture(scope="module")
def harness(tmp_path_factory):
    out = str(tmp_path_factory.mktemp("harness") / "map_harness")
    host = os.path.join(PKG, "host")
    cmd = ["g++", "-O1", "-std=c++17", "-I", os.path.join(ROOT, "include"), "-I", host, "-o", out,
           os.path.join(ROOT, "tests", "host", "map_harness.cpp")] + \
          [os.path.join(host, f) for f in ("bm_classes.cpp", "bm_io.cpp", "bm_locator.cpp")] + \
          ["-L", PKG, "-lbucketmap_b200", "-Wl,-rpath," + PKG, "-pthread"]
    subprocess.check_call(cmd)
    return out


@pytest.mark.parametrize("case", ["full300", "long800", "multirec", "short150"])
def test_mapper_map_lists_match_oracle(oracle, harness, tmp_path, case):
    from bucketmap_b200 import datagen
    g = Golden(case)
    d = str(tmp_path)
    fa, fq = g.write_inputs(d)
    kw = g.params_kw()
    P = oracle.params(g.L, **kw)
    q = kw.get("q", 9)
    oracle.cli_index(fa, g.L, kw.get("r", 300), q, g.frac, g.x, g.y, d, "idx")
    p = subprocess.run([harness, fa, str(g.L), str(P.read_len), str(P.k), str(q), str(P.mapper_samples), str(P.mapper_fault),
                        repr(float(P.distinguishability)), str(P.avg_base_quality), d, "idx", fq],
                       stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, timeout=300)
    assert p.returncode == 0, p.stderr[-2000:]
    lines = p.stdout.splitlines()
    # expected lists from the oracle's per-segment candidate lists (q_gram_mapper.h:527-533)
    NB, b_start, b_len, _ = datagen.bucketize(g.rec_len, g.L, kw.get("r", 300))
    k2i, n_rows = oracle.fracminhash(q, g.x, g.y, g.frac)
    rows = oracle.index_build(g.packed, b_start, b_len, NB, q, k2i, n_rows)
    segs = oracle.segments(P, g.read_off)
    coff, cand, stats = oracle.map_segments(P, NB, k2i, rows, g.seq, g.qual, g.read_off, segs)
    want = {"O": {}, "R": {}}
    for s in range(len(segs[0])):
        item = "%d:%d" % (segs[0][s], segs[1][s])
        for c in cand[coff[2 * s]:coff[2 * s + 1]]:
            want["O"].setdefault(int(c), []).append(item)
        for c in cand[coff[2 * s + 1]:coff[2 * s + 2]]:
            want["R"].setdefault(int(c), []).append(item)
    assert lines[0] == "num_records %d" % (len(g.read_off) - 1)
    assert lines[1] == "buckets %d %d" % (NB, NB)                        # NUM_BUCKETS vectors, phantom buckets included
    got = {"O": {}, "R": {}}
    for ln in lines[2:-1]:
        f = ln.split()
        got[f[0]][int(f[1])] = f[2:]
    assert got == want
    assert sum(len(v) for v in want["O"].values()) > 0
    assert lines[-1] == "after_reset 0"
    assert "The q-gram index is empty. Cannot accept query." in p.stderr    # map() after reset (q_gram_mapper.h:389-393)
    # the log counter counts reads, not segments (long800: five segments per read)
    assert "at least one candidate bucket: %d " % int(stats[0]) in p.stderr
