"""GPU: parity AT SCALE against the reference binary itself (oracle/_ref = the reference's own main.cpp and headers).

BASELINE.json's target is "bit-exact SAM output versus reference bucketmap on a synthetic 1.7 Gbp genome": the goldens under
tests/golden pin the oracle on small inputs; here the reference binaries run on the full-size configurations -- as one
single-threaded process per host core, each on its own shard of the reads -- and the product command line
(bucket-map_b200/bin/bucketmap[_align]) maps the same reads in one run.  The product's SAM must equal the reference's
header followed by the shards' records, BYTE FOR BYTE.

  C2   1.7 Gbp / L=65536, 300 bp reads, -f 1 (the headline) and the default -f 0.25      (>= 100 k reads each)
  C5   the same genome, high-error reads through bucketmap_align (POS, MAPQ, CIGAR)        (>= 20 k reads)
  C1   4.6 Mbp / L=65536, 100 k x 150 bp reads, default flags, exactly as BASELINE states it
  C4   1.7 Gbp / L=262144 at -k 9: no q-gram passes -d 0.5 -- no record, in the reference as here

Each reference process draws its reads from its own span of the genome, so that it builds the per-bucket k-mer index of a
few hundred buckets instead of all 25,940 (8 ms each); that changes nothing about what is compared."""
import os
import shutil
import subprocess
import tempfile

import pytest

import bm_ref
from bucketmap_b200 import datagen

pytestmark = [pytest.mark.gpu, pytest.mark.timeout(1500)]

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, "bucket-map_b200", "bin")
X, Y = 40503, 7919


def _procs():
    try:
        import psutil
        by_mem = int(psutil.virtual_memory().available // (6 << 30))
    except Exception:
        by_mem = 8
    return max(1, min(os.cpu_count() or 1, 16, by_mem))


def _tmp():
    base = "/dev/shm" if os.path.isdir("/dev/shm") else None
    return tempfile.mkdtemp(prefix="bm_scale_", dir=base)


class Genome:
    def __init__(self, n_bases, seed):
        self.n_bases = n_bases
        self.dir = _tmp()
        self.packed = datagen.random_genome_packed(n_bases, seed)
        self.fa = os.path.join(self.dir, "genome.fa")
        datagen.write_fasta_fast(self.fa, self.packed, n_bases, "chr1")
        self.index_dirs = {}

    def index(self, L, q, frac):
        """index files written by the product's `bucketmap -x` (byte-identical to the reference's: tests/golden)"""
        key = (L, q, frac)
        if key not in self.index_dirs:
            d = os.path.join(self.dir, "idx_%d_%d_%g" % key)
            os.makedirs(d)
            p = subprocess.run([os.path.join(BIN, "bucketmap"), "-x", "-i", "idx", "-g", self.fa, "-L", str(L), "-k", str(q),
                                "-f", repr(float(frac)), "--seed-x", str(X), "--seed-y", str(Y)], cwd=d,
                               stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, timeout=900)
            assert p.returncode == 0, p.stderr[-2000:]
            self.index_dirs[key] = d
        return self.index_dirs[key]

    def close(self):
        shutil.rmtree(self.dir, ignore_errors=True)


@pytest.fixture(scope="module")
def genome_1g7():
    g = Genome(1_700_000_000, 3)
    yield g
    g.close()


def _write_fastq(path, seq, qual, off, first_name):
    """FASTQ with numeric read names first_name, first_name+1, ... (so the shards' records concatenate to the whole)"""
    n = len(off) - 1
    with open(path, "wb") as f:
        chunks = []
        for i in range(n):
            a, b = int(off[i]), int(off[i + 1])
            chunks.append(b"@%d\n" % (first_name + i) + seq[a:b].tobytes() + b"\n+\n" + qual[a:b].tobytes() + b"\n")
            if len(chunks) >= 4096:
                f.write(b"".join(chunks))
                chunks = []
        f.write(b"".join(chunks))


def _run_case(g, NB, L, q, frac, read_len, sub, indel, n_per_proc, seed, align=False, spans=True):
    """reference processes on shards + one product run on all reads; returns (product sam, expected sam, records, reads)"""
    if not bm_ref.available(NB, L, align):
        pytest.skip("oracle/_ref binary for NB=%d L=%d is not built (make -C oracle ref)" % (NB, L))
    procs = _procs()
    idx = g.index(L, q, frac)
    work = tempfile.mkdtemp(prefix="case_", dir=g.dir)
    span = max(int(g.n_bases * n_per_proc / 1_000_000), 4 * L) if spans else g.n_bases
    span = min(span - span % 4, g.n_bases)
    if spans:
        procs = max(1, min(procs, g.n_bases // span))
    env = dict(os.environ, BM_GENOME=g.fa, BM_RAND=f"{X - 1},{Y}")
    runs = []
    all_fq = os.path.join(work, "all.fq")
    with open(all_fq, "wb") as allf:
        for pi in range(procs):
            d = os.path.join(work, f"p{pi}")
            os.makedirs(d)
            for ext in ("qgram", "kmers_index", "bucket_id"):
                os.symlink(os.path.join(idx, "idx." + ext), os.path.join(d, "idx." + ext))
            lo = ((g.n_bases - span) // max(1, procs - 1) * pi if procs > 1 else 0) if spans else 0
            lo -= lo % 4
            seq, qual, off = datagen.simulate_reads(g.packed[lo // 4:], span, n_per_proc, read_len, sub, indel, indel,
                                                    seed + pi)[:3]
            fq = os.path.join(d, "reads.fq")
            _write_fastq(fq, seq, qual, off, pi * n_per_proc)
            allf.write(open(fq, "rb").read())
            cmd = [bm_ref.binary(NB, L, align), "-i", "idx", "-q", fq, "-o", os.path.join(d, "out.sam"), "-f", repr(float(frac)),
                   "-k", str(q)]
            runs.append(subprocess.Popen(cmd, cwd=d, env=env, stdout=subprocess.DEVNULL, stderr=subprocess.PIPE, text=True))
    # the product maps all reads in one run while the reference processes work
    pd = os.path.join(work, "product")
    os.makedirs(pd)
    for ext in ("qgram", "kmers_index", "bucket_id"):
        os.symlink(os.path.join(idx, "idx." + ext), os.path.join(pd, "idx." + ext))
    exe = os.path.join(BIN, "bucketmap_align" if align else "bucketmap")
    pp = subprocess.run([exe, "-i", "idx", "-q", all_fq, "-o", "out.sam", "-g", g.fa, "-L", str(L), "-k", str(q),
                         "-f", repr(float(frac)), "--seed-x", str(X), "--seed-y", str(Y)], cwd=pd,
                        stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, timeout=1200)
    logs = [r.communicate()[1] for r in runs]
    assert pp.returncode == 0, pp.stderr[-3000:]
    for r, lg in zip(runs, logs):
        assert r.returncode == 0, lg[-2000:]
    header, body = None, []
    for pi in range(procs):
        text = open(os.path.join(work, f"p{pi}", "out.sam")).read()
        lines = text.splitlines(keepends=True)
        h = [ln for ln in lines if ln.startswith("@")]
        if header is None:
            header = h
        assert h == header
        body += [ln for ln in lines if not ln.startswith("@")]
    want = "".join(header) + "".join(body)
    got = open(os.path.join(pd, "out.sam")).read()
    shutil.rmtree(work, ignore_errors=True)
    return got, want, len(body), procs * n_per_proc


def _assert_same(got, want):
    if got != want:
        gl, wl = got.splitlines(), want.splitlines()
        assert len(gl) == len(wl), (len(gl), len(wl))
        for i, (a, b) in enumerate(zip(gl, wl)):
            assert a == b, "first difference at line %d" % i
    assert got == want


@pytest.mark.parametrize("frac", [1.0, 0.25])
def test_c2_sam_equals_reference_binary(genome_1g7, frac):
    got, want, n_rec, n_reads = _run_case(genome_1g7, 25940, 65536, 9, frac, 300, 0.002, 0.00025, 6400, 9000)
    assert n_reads >= 100_000 or _procs() < 16
    assert n_rec > 0.95 * n_reads
    _assert_same(got, want)


def test_c5_align_sam_equals_reference_binary(genome_1g7):
    got, want, n_rec, n_reads = _run_case(genome_1g7, 25940, 65536, 9, 1.0, 300, 0.01, 0.001, 1300, 9100, align=True)
    assert n_reads >= 20_000 or _procs() < 16
    assert n_rec > 0.9 * n_reads
    _assert_same(got, want)


def test_c4_262144_k9_maps_nothing_like_the_reference(genome_1g7):
    """a 262,144-base bucket holds 63 % of all 9-mers: no q-gram is distinguishable at -d 0.5, no read has a candidate"""
    got, want, n_rec, n_reads = _run_case(genome_1g7, 6485, 262144, 9, 1.0, 300, 0.002, 0.00025, 500, 9200)
    assert n_rec == 0
    _assert_same(got, want)


def test_c1_as_baseline_states_it():
    """4.6 Mbp reference, L=65536, 100 k x 150 bp reads, default flags (-f 0.25, -r 300)"""
    g = Genome(4_600_000, 1)
    try:
        got, want, n_rec, n_reads = _run_case(g, 71, 65536, 9, 0.25, 150, 0.002, 0.00025, 6250, 2, spans=False)
        assert n_reads == 100_000 or _procs() < 16
        assert n_rec > 0.9 * n_reads
        _assert_same(got, want)
    finally:
        g.close()
