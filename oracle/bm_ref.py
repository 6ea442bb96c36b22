"""Driver of oracle/_ref -- the reference's own bucketmap / bucketmap_align built from /root/reference against the
shim headers (oracle/Makefile).  TEST INFRASTRUCTURE ONLY: allowed importers are tests/, tests/golden/make_golden.py
and bench.py's cpu_baseline / --impl reference legs.  Nothing under bucket-map_b200/ imports this.

The binaries are per (NB, L) because the reference bakes both into the build; building needs /root/reference (this
container), running needs only the prebuilt file (it travels to the GPU box with the snapshot)."""
import os
import re
import subprocess

_HERE = os.path.dirname(os.path.abspath(__file__))
REF_DIR = os.path.join(_HERE, "_ref")
REFERENCE_SRC = "/root/reference/bucket_map"


def binary(NB, L, align=False):
    return os.path.join(REF_DIR, f"{'bucketmapalign' if align else 'bucketmap'}_{NB}_{L}")


def available(NB, L, align=False):
    return os.path.exists(binary(NB, L, align))


def can_build():
    return os.path.isdir(REFERENCE_SRC)


def build(NB, L):
    """compile both flavours for (NB, L); no-op when they exist and are newer than their sources"""
    if not can_build():
        raise RuntimeError("the reference sources (/root/reference) are not present on this machine")
    subprocess.check_call(["make", "-s", "-C", _HERE, "ref", f"NB={NB}", f"L={L}"])
    return binary(NB, L, False), binary(NB, L, True)


def _flags(kw):
    """CLI flags of main.cpp:41-124 from keyword names (k q-gram length, l query seed, r, s, d, b, e, n, p, u, f)"""
    out = []
    for key in ("k", "l", "r", "s", "d", "b", "e", "n", "p", "u", "f"):
        if key in kw and kw[key] is not None:
            out += [f"-{key}", repr(kw[key]) if isinstance(kw[key], float) else str(kw[key])]
    return out


def _run(cmd, cwd, genome, x, y, timeout):
    env = dict(os.environ, BM_GENOME=genome, BM_RAND=f"{x - 1},{y}")   # x = rand() % (p-1) + 1, y = rand() % p
    p = subprocess.run(cmd, cwd=cwd, env=env, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True,
                       timeout=timeout)
    if p.returncode != 0:
        raise RuntimeError(f"{' '.join(cmd)} exited {p.returncode}\n{p.stderr[-2000:]}")
    return p.stderr


def index(NB, L, workdir, genome, indicator, x, y, timeout=3600, **kw):
    """bucketmap -x -i <indicator>  (writes <workdir>/<indicator>.{qgram,kmers_index,bucket_id})"""
    return _run([binary(NB, L), "-x", "-i", indicator] + _flags(kw), workdir, genome, x, y, timeout)


def map_reads(NB, L, workdir, genome, indicator, fastq, sam, x, y, align=False, timeout=7200, **kw):
    """bucketmap[_align] -i <indicator> -q <fastq> -o <sam>; returns the [BENCHMARK] phase timers parsed from stderr"""
    if os.path.exists(sam):
        os.remove(sam)                     # sharg's output validator refuses an existing file (main.cpp:68)
    log = _run([binary(NB, L, align), "-i", indicator, "-q", fastq, "-o", sam] + _flags(kw), workdir, genome, x, y,
               timeout)
    return parse_log(log)


_PATTERNS = {
    "load_s": r"Elapsed time for loading index files: ([0-9.eE+-]+) s",
    "map_s": r"Elapsed time for bucket mapping: ([0-9.eE+-]+) s",
    "with_candidate": r"at least one candidate bucket: (\d+)",
    "kmer_index_s": r"building k-mer index for each bucket: ([0-9.eE+-]+) s",
    "find_offset_s": r"finding exact location of the sequences: ([0-9.eE+-]+) s",
    "locations": r"Total mapped locations: (\d+)",
    "output_s": r"alignment verification and output: ([0-9.eE+-]+) s",
}


def parse_log(log):
    out = {"log": log}
    for k, pat in _PATTERNS.items():
        m = re.search(pat, log)
        if m:
            out[k] = float(m.group(1)) if k.endswith("_s") else int(m.group(1))
    return out
