"""GPU: the drop-in command line (bucket-map_b200/bin/bucketmap, bucketmap_align: C++ host over the C-ABI) against the
committed outputs of the reference's own bucketmap on the same inputs (tests/golden): index files byte-exact, SAM text
byte-exact."""
import os
import subprocess

import pytest

from golden_util import ALIGN_CASES, CASES, Golden, sam_records, sha256_file

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, "bucket-map_b200", "bin")


def _flags(g):
    out = []
    for k, v in g.flags.items():
        out += [f"-{k}", repr(v) if isinstance(v, float) else str(v)]
    return out


def _run(exe, args, cwd, env=None):
    p = subprocess.run([os.path.join(BIN, exe)] + args, cwd=cwd, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True,
                       timeout=600, env=dict(os.environ, **env) if env else None)
    assert p.returncode == 0, p.stderr[-3000:]
    return p.stderr


def _index(g, d):
    fa, fq = g.write_inputs(d)
    log = _run("bucketmap", ["-x", "-i", "idx", "-g", fa, "-L", str(g.L), "--seed-x", str(g.x), "--seed-y", str(g.y)]
               + _flags(g), d)
    return fa, fq, log


def _assert_sam_equal(got, want):
    if got != want:
        gh, gr = sam_records(got)
        wh, wr = sam_records(want)
        assert gh == wh
        assert len(gr) == len(wr)
        for a, b in zip(gr, wr):
            assert a == b
    assert got == want


@pytest.mark.parametrize("case", CASES)
def test_cli_index_files_match_reference(tmp_path, case):
    g = Golden(case)
    d = str(tmp_path)
    _index(g, d)
    assert sha256_file(tmp_path / "idx.qgram") == g.meta["qgram_sha256"]
    assert sha256_file(tmp_path / "idx.kmers_index") == g.meta["kmers_index_sha256"]
    assert open(tmp_path / "idx.bucket_id").read() == g.meta["bucket_id"]


@pytest.mark.parametrize("case", CASES)
def test_cli_sam_matches_reference(tmp_path, case):
    g = Golden(case)
    d = str(tmp_path)
    fa, fq, _ = _index(g, d)
    log = _run("bucketmap", ["-i", "idx", "-q", fq, "-o", "out.sam", "-g", fa, "-L", str(g.L)] + _flags(g), d)
    assert "already exists" in log                      # index() refuses to overwrite and the run goes on (note 14)
    _assert_sam_equal(open(tmp_path / "out.sam").read(), g.sam(False))
    assert f"Total mapped locations: {g.meta['ref_counters']['locations']} " in log


@pytest.mark.parametrize("case", ALIGN_CASES)
def test_cli_align_sam_matches_reference(tmp_path, case):
    g = Golden(case)
    d = str(tmp_path)
    fa, fq, _ = _index(g, d)
    _run("bucketmap_align", ["-i", "idx", "-q", fq, "-o", "out.sam", "-g", fa, "-L", str(g.L)] + _flags(g), d)
    _assert_sam_equal(open(tmp_path / "out.sam").read(), g.sam(True))


@pytest.mark.parametrize("case", ["repeats", "lowqual", "long800"])
def test_cli_table_free_kernels_give_the_same_sam(tmp_path, case):
    """$BM_KMER_TABLE=0 / $BM_LOCATOR_INDEX=0: q-gram row gather and bucket text scan instead of the derived tables"""
    g = Golden(case)
    d = str(tmp_path)
    fa, fq, _ = _index(g, d)
    _run("bucketmap", ["-i", "idx", "-q", fq, "-o", "a.sam", "-g", fa, "-L", str(g.L)] + _flags(g), d,
         env={"BM_KMER_TABLE": "0", "BM_LOCATOR_INDEX": "0"})
    _assert_sam_equal(open(tmp_path / "a.sam").read(), g.sam(False))


def test_cli_small_batches_give_the_same_sam(tmp_path):
    """--batch-reads only changes how the FASTQ is cut into device batches"""
    g = Golden("repeats")
    d = str(tmp_path)
    fa, fq, _ = _index(g, d)
    _run("bucketmap", ["-i", "idx", "-q", fq, "-o", "a.sam", "-g", fa, "-L", str(g.L), "--batch-reads", "37"] + _flags(g), d)
    _assert_sam_equal(open(tmp_path / "a.sam").read(), g.sam(False))


def test_cli_host_threads_do_not_change_the_sam(tmp_path):
    """--threads only changes how many host threads pack the FASTA / FASTQ and render the SAM"""
    g = Golden("multirec")
    d = str(tmp_path)
    fa, fq, _ = _index(g, d)
    for t in ("1", "3", "16"):
        _run("bucketmap", ["-i", "idx", "-q", fq, "-o", f"t{t}.sam", "-g", fa, "-L", str(g.L), "--threads", t,
                           "--batch-reads", "101"] + _flags(g), d)
        _assert_sam_equal(open(tmp_path / f"t{t}.sam").read(), g.sam(False))


def _n_gpus():
    try:
        import torch
        return torch.cuda.device_count()
    except Exception:
        return 1


@pytest.mark.skipif(_n_gpus() < 2, reason="needs two GPUs")
@pytest.mark.parametrize("case,align", [("repeats", False), ("full300", True)])
def test_cli_several_devices_give_the_same_sam(tmp_path, case, align):
    """--devices spreads the batches of one FASTQ over the GPUs (index replicated over NCCL); records leave in read order"""
    g = Golden(case)
    d = str(tmp_path)
    fa, fq, _ = _index(g, d)
    n = min(_n_gpus(), 4)
    _run("bucketmap_align" if align else "bucketmap",
         ["-i", "idx", "-q", fq, "-o", "m.sam", "-g", fa, "-L", str(g.L), "--devices", "0-%d" % (n - 1), "--batch-reads", "23"]
         + _flags(g), d)
    _assert_sam_equal(open(tmp_path / "m.sam").read(), g.sam(align))


def test_cli_builds_index_when_missing_then_maps(tmp_path):
    """`bucketmap -i -q -o` without a prior -x builds the index first (locator::initialize, locator.h:22-36)"""
    g = Golden("full300")
    d = str(tmp_path)
    fa, fq = g.write_inputs(d)
    _run("bucketmap", ["-i", "idx", "-q", fq, "-o", "out.sam", "-g", fa, "-L", str(g.L), "--seed-x", str(g.x),
                       "--seed-y", str(g.y)] + _flags(g), d)
    assert sha256_file(tmp_path / "idx.qgram") == g.meta["qgram_sha256"]
    _assert_sam_equal(open(tmp_path / "out.sam").read(), g.sam(False))


def test_cli_legacy_index_without_kmers_index_maps_nothing_like_the_reference(tmp_path):
    """An index written before FracMinHash existed has no .kmers_index file.  The reference then reads no table and zero
    rows (q_gram_mapper.h:331-357), maps nothing and writes a SAM with the header only, exit code 0.
    tests/golden/full300.legacy_noindex.sam is what oracle/_ref/bucketmap_49_8192 wrote for the full300 inputs after
    `rm idx.kmers_index` (observed in the build container)."""
    g = Golden("full300")
    d = str(tmp_path)
    fa, fq, _ = _index(g, d)
    os.remove(tmp_path / "idx.kmers_index")
    log = _run("bucketmap", ["-i", "idx", "-q", fq, "-o", "out.sam", "-g", fa, "-L", str(g.L)] + _flags(g), d)
    want = open(os.path.join(ROOT, "tests", "golden", "full300.legacy_noindex.sam")).read()
    assert open(tmp_path / "out.sam").read() == want
    assert "Total mapped locations: 0 " in log


def test_cli_argument_errors(tmp_path):
    d = str(tmp_path)
    p = subprocess.run([os.path.join(BIN, "bucketmap"), "-x"], cwd=d, stderr=subprocess.PIPE, text=True)
    assert p.returncode == 255 and "[ERROR]" in p.stderr          # main.cpp:151-155 returns -1
    open(tmp_path / "r.fq", "w").write("@r\nACGT\n+\nEEEE\n")
    open(tmp_path / "o.sam", "w").write("")
    p = subprocess.run([os.path.join(BIN, "bucketmap"), "-i", "x", "-q", "r.fq", "-o", "o.sam"], cwd=d,
                       stderr=subprocess.PIPE, text=True)
    assert p.returncode == 255 and "already exists" in p.stderr   # sharg output_file_validator (main.cpp:68)
