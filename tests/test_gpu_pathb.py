"""GPU: INTEGRATION.md path B -- the reference's own main.cpp with the three classes swapped for bm::bucket_hash_indexer /
bm::q_gram_mapper / bm::bucket_locator (oracle/pathb.py applies the patch and builds it here, where /root/reference is)
-- reproduces the reference's index files and SAM on the goldens, byte for byte."""
import os
import subprocess

import pytest

import pathb
from golden_util import Golden, sha256_file

pytestmark = pytest.mark.gpu

CASES = ["full300", "repeats", "long800", "higherr", "short150"]


def _flags(g):
    out = []
    for k, v in g.flags.items():
        out += [f"-{k}", repr(v) if isinstance(v, float) else str(v)]
    return out


def _run(exe, args, cwd, g, fa):
    env = dict(os.environ, BM_GENOME=fa, BM_RAND=f"{g.x - 1},{g.y}")
    p = subprocess.run([exe] + args, cwd=cwd, env=env, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, timeout=600)
    assert p.returncode == 0, p.stderr[-3000:]
    return p.stderr


@pytest.mark.parametrize("case", CASES)
def test_patched_reference_main_reproduces_the_goldens(tmp_path, case):
    g = Golden(case)
    if not pathb.available(g.NB, g.L):
        pytest.skip("oracle/_ref/pathb_* not built (python oracle/pathb.py, needs /root/reference)")
    d = str(tmp_path)
    fa, fq = g.write_inputs(d)
    _run(pathb.binary(g.NB, g.L), ["-x", "-i", "idx"] + _flags(g), d, g, fa)
    assert sha256_file(tmp_path / "idx.qgram") == g.meta["qgram_sha256"]
    assert sha256_file(tmp_path / "idx.kmers_index") == g.meta["kmers_index_sha256"]
    assert open(tmp_path / "idx.bucket_id").read() == g.meta["bucket_id"]
    log = _run(pathb.binary(g.NB, g.L), ["-i", "idx", "-q", fq, "-o", "out.sam"] + _flags(g), d, g, fa)
    assert open(tmp_path / "out.sam").read() == g.sam(False)
    assert "Not using Smith-Waterman" in log                   # the reference main's own banner (main.cpp:160-163)
    if os.path.exists(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", case + ".align.sam.gz")):
        _run(pathb.binary(g.NB, g.L, True), ["-i", "idx", "-q", fq, "-o", "aln.sam"] + _flags(g), d, g, fa)
        assert open(tmp_path / "aln.sam").read() == g.sam(True)
