#!/usr/bin/env python
"""bench.py -- reads/sec mapped through the candidate-bucket hot path (BASELINE.json metric) on N B200s.

A step is one pass of the hot path (mapper K2f = k-mer filter + sampler + coarse vote + exact recount, locator K3i + gather,
best-location merge K3b; for c5 the bucketmap_align verification K4 instead of the merge) over one batch of synthetic reads.
Default workload at any N: BASELINE `configs[1]` ("synthetic 1.7 Gbp reference, BUCKET_LEN=65536, 1M x 300bp reads, 0.2%
sub, 0.025% ins/del") per GPU -- reads are sharded, the index is replicated (weak scaling).  --config picks the other
BASELINE configurations (c1, c3, c4-16384, c4-262144[-k10], c5) and the repeat-rich C2' (c2r); --frac the FracMinHash
fraction (-f; the reference's default is 0.25, the headline line is quoted on -f 1).

  value     whole-job reads/s with the packed read batch resident in HBM (CUDA events on the library's stream, max over ranks)
  e2e       the same through bm_map_locate_packed() with pinned HOST buffers in the form the FASTQ parser of the C++ host
            produces (2-bit base planes; the k-mer quality plane only when a read has a base below -b): H2D of the batch
            and D2H of the location records inside the timed region.  e2e_ascii: the same through bm_map_locate() with the
            FASTQ's ASCII bases + qualities (6.6x the bytes; converted on the device).
  roofline  the dominant kernel of the step (k_map_fused; k_align for c5): algorithmic bytes = the bytes its loads ask for,
            counted in the kernel; time = CUDA events around each launch inside the timed region; traffic = DRAM bytes of
            the committed ncu capture of THIS build of the kernel (profiles/roofline_traffic.json carries the hash of the
            kernel sources it was taken from; a stale capture is reported as null).  roofline_qgram_path: the same for
            k_map_vote, the full-width q-gram row gather used when the k-mer tables are off or do not fit.
  cpu_baseline / --impl reference: the reference's CPU implementation of the path on the host cores, on a bounded sample of
            the same workload: oracle/_ref (the reference's own main.cpp and headers, built from /root/reference against
            the test-only seqan3 shim; "reference") as one single-threaded process per core, else the plain C++
            restatement under oracle/ with all host threads ("port").
  parity    (N=1, with the cpu_baseline leg) the SAM files the reference processes wrote for their sample reads, compared
            field by field with the records bm_map_locate returns for the same reads.
  N > 1     reads sharded over the ranks, index replicated (ncclBroadcast from rank 0, bm_bcast_index); each step ends with
            the gather of the location records into rank 0's device memory (bm_gather_results: peer-memory stores over
            NVLink, packed by rank 0, overlapping the next step); CUDA events as at N=1, max over ranks.  gather_verified:
            rank 0 re-maps another rank's shard locally and compares it with the gathered records.
"""
import argparse
import json
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

CONFIGS = {
    # genome bases, genome seed, bucket_len, index seed q, read_len, read seed, sub, indel, default reads per GPU per step
    "c1": dict(n_bases=4_600_000, gseed=1, L=65536, q=9, read_len=150, rseed=2, sub=0.002, indel=0.00025, reads=100_000,
               desc="4.6 Mbp synthetic reference, L=65536, 100k x 150bp reads (0.2% sub, 0.025% ins/del)"),
    "c2": dict(n_bases=1_700_000_000, gseed=3, L=65536, q=9, read_len=300, rseed=4, sub=0.002, indel=0.00025, reads=1_000_000,
               desc="1.7 Gbp synthetic reference, L=65536, 300bp reads (0.2% sub, 0.025% ins/del)"),
    "c2r": dict(n_bases=1_700_000_000, gseed=3, L=65536, q=9, read_len=300, rseed=4, sub=0.002, indel=0.00025, reads=1_000_000,
                repeats=(10_000, 8_500),
                desc="C2': the 1.7 Gbp reference with 5% of it overwritten by 8,500 duplicated 10 kbp segments, 300bp reads"),
    "c2x": dict(n_bases=1_700_000_000, gseed=3, L=65536, q=9, read_len=300, rseed=4, sub=0.002, indel=0.00025, reads=1_000_000,
                family=(1_000, 34_000, 0.03),
                desc="repeat-family stress: the 1.7 Gbp reference with 2% of it = 34,000 copies of one 1 kbp element (3% "
                     "divergence, a copy in almost every bucket), 300bp reads"),
    "c3": dict(n_bases=3_100_000_000, gseed=5, L=65536, q=9, read_len=150, rseed=6, sub=0.002, indel=0.00025, reads=10_000_000,
               strong=True, desc="3.1 Gbp synthetic reference, L=65536, 10M x 150bp reads sharded over the GPUs"),
    "c4-16384": dict(n_bases=1_700_000_000, gseed=3, L=16384, q=9, read_len=300, rseed=4, sub=0.002, indel=0.00025,
                     reads=1_000_000, desc="bucket-length sweep: 1.7 Gbp reference, L=16384, 300bp reads"),
    "c4-262144": dict(n_bases=1_700_000_000, gseed=3, L=262144, q=9, read_len=300, rseed=4, sub=0.002, indel=0.00025,
                      reads=1_000_000, desc="bucket-length sweep: 1.7 Gbp reference, L=262144, 300bp reads (-k 9: no q-gram "
                                            "passes -d 0.5, nothing maps, in the reference as here)"),
    "c4-262144-k10": dict(n_bases=1_700_000_000, gseed=3, L=262144, q=10, read_len=300, rseed=4, sub=0.002, indel=0.00025,
                          reads=1_000_000, desc="bucket-length sweep: 1.7 Gbp reference, L=262144, -k 10, 300bp reads"),
    "c5": dict(n_bases=1_700_000_000, gseed=3, L=65536, q=9, read_len=300, rseed=7, sub=0.01, indel=0.001, reads=1_000_000,
               align=True, desc="bucketmap_align path: 1.7 Gbp reference, L=65536, 300bp high-error reads (1% sub, 0.2% "
                                "indel), in-bucket bit-parallel verification with CIGAR"),
}
FMH_X, FMH_Y = 40503, 7919


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--config", default="c2", choices=sorted(CONFIGS))
    ap.add_argument("--reads", type=int, default=0, help="reads per GPU per step (0: the configuration's own)")
    ap.add_argument("--frac", type=float, default=1.0, help="FracMinHash fraction (-f): 1 = the headline, 0.25 = the reference's default")
    ap.add_argument("--ftable", type=int, default=1, choices=[0, 1], help="full-resolution k-mer table for the exact recount")
    ap.add_argument("--e2e-steps", type=int, default=0, help="timed end-to-end steps (0: as --steps)")
    ap.add_argument("--e2e-lanes", type=int, default=2, choices=[1, 2],
                    help="host threads (calls in flight) per GPU in the end-to-end steps: 2 = the context and a lane")
    ap.add_argument("--cpu-sample", type=int, default=0, help="reads in the CPU baseline sample (0 = auto)")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--lanes", type=int, default=0,
                    help="bm_map_locate runs batches of at least this many reads on two pipeline lanes (0: library default)")
    ap.add_argument("--lidx", type=int, default=1, choices=[0, 1],
                    help="1: locator searches the genome k-mer position index; 0: scans each candidate bucket's text")
    ap.add_argument("--ktable", type=int, default=1, choices=[0, 1, 8, 16, 32],
                    help="1: mapper votes over the coarse k-mer table (derived on the device when 4^k rows fit; 16 "
                         "buckets per coarse bit; 8/16/32 pick the group size); 0: gathers the k-q+1 q-gram rows per "
                         "sample over the full row width")
    return ap.parse_args()


class ClockSampler(threading.Thread):
    """samples SM clock and throttle reasons of one GPU while the timed region runs (pynvml)"""

    def __init__(self, index, period=0.05):
        super().__init__(daemon=True)
        self.index, self.period = index, period
        self.sm, self.reasons, self.sm_max = [], set(), None
        self._halt = threading.Event()
        self.ok = False
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.sm_max = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
            self.ok = True
        except Exception:
            self.ok = False

    def run(self):
        if not self.ok:
            return
        nv = self.nv
        names = {
            getattr(nv, "nvmlClocksThrottleReasonHwSlowdown", 0x8): "hw_slowdown",
            getattr(nv, "nvmlClocksThrottleReasonHwThermalSlowdown", 0x40): "hw_thermal_slowdown",
            getattr(nv, "nvmlClocksThrottleReasonSwThermalSlowdown", 0x20): "sw_thermal_slowdown",
            getattr(nv, "nvmlClocksThrottleReasonSwPowerCap", 0x4): "sw_power_cap",
        }
        while not self._halt.is_set():
            try:
                self.sm.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for bit, name in names.items():
                    if r & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            self._halt.wait(self.period)

    def stop(self):
        self._halt.set()
        if self.is_alive():
            self.join()
        return {"sm_mhz": float(np.median(self.sm)) if self.sm else None, "sm_max_mhz": self.sm_max,
                "reasons": sorted(self.reasons)}


def make_workload(cfg):
    """genome (packed) + buckets of a configuration"""
    from bucketmap_b200 import datagen
    packed = datagen.random_genome_packed(cfg["n_bases"], cfg["gseed"])
    if cfg.get("repeats"):
        seg, copies = cfg["repeats"]
        packed = datagen.add_repeats_packed(packed, cfg["n_bases"], seg, copies, cfg["gseed"] + 100)
    if cfg.get("family"):
        unit, copies, div = cfg["family"]
        packed = datagen.add_family_packed(packed, cfg["n_bases"], unit, copies, div, cfg["gseed"] + 200)
    NB, b_start, b_len, _ = datagen.bucketize(np.array([cfg["n_bases"]], np.uint64), cfg["L"], 300)
    return packed, NB, b_start, b_len


def simulate(cfg, packed, n_reads, seed, n_span=None):
    from bucketmap_b200 import datagen
    return datagen.simulate_reads(packed, n_span or cfg["n_bases"], n_reads, cfg["read_len"], cfg["sub"],
                                  cfg["indel"], cfg["indel"], seed)[:3]


def accuracy(loff, locs, L, start, rev, tolerance=10):
    """the analyzer's metrics (benchmark/sam_file_analyzer.cpp:250-358, tolerance 10 as its main() sets) of one batch's
    location records against the simulator's ground truth; one reference sequence, so position = bucket * L + offset"""
    cnt = np.diff(loff.astype(np.int64))
    read_of = np.repeat(np.arange(len(cnt)), cnt)
    pos = locs["bucket"].astype(np.int64) * L + locs["offset"]
    good = (np.abs(pos - start[read_of]) <= tolerance) & ((locs["is_original"] == 0) == rev[read_of])
    correct = np.zeros(len(cnt), bool)
    correct[read_of[good]] = True
    single = cnt <= 1                     # the analyzer's per-read counts leave out reads with several records
    n = len(cnt)
    return {"reads": int(n), "mapped_pct": 100.0 * int(((cnt > 0) & single).sum()) / n,
            "correct_pct": 100.0 * int((correct & single).sum()) / n,
            "reads_with_several_locations": int((~single).sum()),
            "correct_incl_several_pct": 100.0 * int(correct.sum()) / n,
            "locations": int(len(locs)), "precision_pct": 100.0 * int(good.sum()) / max(1, len(locs)),
            "tolerance": tolerance, "definition": "benchmark/sam_file_analyzer.cpp:250-358 on the records of one step "
                                                  "against the read simulator's origin and strand"}


def bind_to_gpu_numa_node(device):
    """run this rank (and first-touch its pinned buffers) on the CPUs next to its GPU: on a two-socket box a rank that
    lands on the far socket pushes every host<->device byte across the socket link"""
    try:
        import pynvml
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(device)
        try:
            pynvml.nvmlDeviceSetCpuAffinity(h)
        except Exception:
            words = pynvml.nvmlDeviceGetCpuAffinity(h, (os.cpu_count() + 63) // 64)
            cpus = [64 * i + b for i, w in enumerate(words) for b in range(64) if (w >> b) & 1]
            if cpus:
                os.sched_setaffinity(0, cpus)
        return sorted(os.sched_getaffinity(0))
    except Exception:
        return None


def measured_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


TRAFFIC_SOURCES = ["bucket-map_b200/csrc/bm_mapfused.cu", "bucket-map_b200/csrc/bm_internal.cuh",
                   "bucket-map_b200/csrc/bm_index.cu"]


def kernel_source_hash():
    import hashlib
    h = hashlib.sha256()
    for f in TRAFFIC_SOURCES:
        h.update(open(os.path.join(ROOT, f), "rb").read())
    return h.hexdigest()[:16]


def ncu_traffic(config_name, frac, n_reads, kernel):
    """per-launch DRAM bytes (read + write) of the dominant kernel from the committed `ncu --set full` capture of this
    workload, scaled from the captured launch to this launch's reads (no reuse across reads: traffic is linear in reads).
    The capture names the hash of the kernel sources it was taken from; if the sources changed since, it is stale and
    None is reported (with the reason) instead of a number that no longer belongs to the kernel."""
    p = os.path.join(ROOT, "profiles", "roofline_traffic.json")
    if not os.path.exists(p):
        return None, "no capture committed"
    try:
        d = json.load(open(p))
        if d.get("workload") != config_name or d.get("kernel") != kernel or abs(d.get("frac", 1.0) - frac) > 1e-9:
            return None, "the committed capture is of another workload / kernel"
        if d.get("source_sha256_16") != kernel_source_hash():
            return None, "stale: the kernel sources changed since the capture (%s)" % d.get("capture")
        per = (d["dram_bytes_read_per_launch"] + d["dram_bytes_write_per_launch"]) / d["captured_reads_per_launch"]
        return per * n_reads, "ncu --set full capture %s of this build (dram__bytes_read.sum + dram__bytes_write.sum, %d reads)" % (
            d.get("capture"), d["captured_reads_per_launch"])
    except Exception as e:      # noqa: BLE001
        return None, "unreadable capture: %s" % e


# ------------------------------------------------------------------------------------------------- CPU (oracle) leg
def cpu_leg(cfg, frac, packed, NB, b_start, b_len, k2i, rows, n_sample, full_reads, threads):
    """times the CPU restatement (map + locate + best-location merge) on a bounded sample of the workload.
    The sample reads start inside the first n_sample/full_reads of the genome, so the number of reads per bucket (the
    amortisation of the reference's per-bucket k-mer index build) matches the full workload."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import bm_oracle
    bm_oracle.build()
    o = bm_oracle.Oracle(threads)
    P = o.params(cfg["L"], q=cfg["q"])
    span = max(int(cfg["n_bases"] * n_sample / full_reads), 4 * cfg["L"])
    span = min(span, cfg["n_bases"])
    seq, qual, off = simulate(cfg, packed, n_sample, cfg["rseed"] + 1000, n_span=span)
    segs = o.segments(P, off)
    ix = o.index_open(P, NB, k2i, rows)
    t0 = time.perf_counter()
    coff, cand, stats = o.map_segments_h(ix, P, seq, qual, off, segs)
    t1 = time.perf_counter()
    loff, locs = o.locate(P, packed, b_start, b_len, seq, qual, off, segs, coff, cand)
    foff, best = o.filter_best(P, off, loff, locs)
    t2 = time.perf_counter()
    o.index_close(ix)
    return {"reads": n_sample, "seconds": t2 - t0, "map_s": t1 - t0, "locate_s": t2 - t1,
            "reads_per_s": n_sample / (t2 - t0), "threads": o.threads, "n_locs": int(len(best)),
            "span": span}


# ----------------------------------------------------------------------------------- CPU (reference itself) leg
def reference_leg(cfg, frac, packed, NB, b_start, k2i, rows, n_per_proc, full_reads, procs, keep=None):
    """times oracle/_ref/bucketmap[align]_<NB>_<L> -- the reference's own main.cpp built against oracle/shim -- on a
    bounded sample of the workload: `procs` independent single-threaded processes (the reference has no threads), each
    mapping its own n_per_proc reads drawn from its own genome span sized so that reads per bucket match the full batch.
    Timed: the reference's own [BENCHMARK] phase timers for bucket mapping + per-bucket k-mer index builds + offset
    finding (+ alignment verification and output for the align flavour); FASTA/index loading is outside, as it is for
    the GPU arm.  The index files are handed over ready-made (they are byte-identical to what `bucketmap -x` writes:
    tests/golden).  keep: a list that receives (seq, qual, off, sam text) of every process for the parity check."""
    import shutil
    import subprocess
    import tempfile
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import bm_ref
    from bucketmap_b200 import datagen
    L, n_bases, align = cfg["L"], cfg["n_bases"], bool(cfg.get("align"))
    work = tempfile.mkdtemp(prefix="bm_ref_bench_")
    try:
        fa = os.path.join(work, "genome.fa")
        datagen.write_fasta_fast(fa, packed, n_bases, "chr1")
        np.ascontiguousarray(rows).tofile(os.path.join(work, "idx.qgram"))
        with open(os.path.join(work, "idx.kmers_index"), "wb") as f:
            f.write(("\n".join(map(str, k2i.tolist())) + "\n").encode())
        with open(os.path.join(work, "idx.bucket_id"), "w") as f:
            f.write("chr1\n" * len(b_start))
        span = max(int(n_bases * n_per_proc / full_reads), 4 * L)
        span -= span % 4
        procs = max(1, min(procs, n_bases // span))
        runs, inputs = [], []
        env = dict(os.environ, BM_GENOME=fa, BM_RAND=f"{FMH_X - 1},{FMH_Y}")
        for pi in range(procs):
            d = os.path.join(work, f"p{pi}")
            os.makedirs(d)
            for ext in ("qgram", "kmers_index", "bucket_id"):
                os.symlink(os.path.join(work, "idx." + ext), os.path.join(d, "idx." + ext))
            lo = (n_bases - span) // max(1, procs - 1) * pi if procs > 1 else 0
            lo -= lo % 4
            seq, qual, off = datagen.simulate_reads(packed[lo // 4:], span, n_per_proc, cfg["read_len"], cfg["sub"],
                                                    cfg["indel"], cfg["indel"], cfg["rseed"] + 1000 + pi)[:3]
            fq = os.path.join(d, "reads.fq")
            datagen.write_fastq_fast(fq, seq, qual, off)
            inputs.append((seq, qual, off))
            cmd = [bm_ref.binary(NB, L, align), "-i", "idx", "-q", fq, "-o", os.path.join(d, "out.sam"), "-f", repr(float(frac)),
                   "-k", str(cfg["q"])]
            runs.append(subprocess.Popen(cmd, cwd=d, env=env, stdout=subprocess.DEVNULL, stderr=subprocess.PIPE, text=True))
        t0 = time.perf_counter()
        logs = [r.communicate()[1] for r in runs]
        wall = time.perf_counter() - t0
        for r, lg in zip(runs, logs):
            if r.returncode != 0:
                raise RuntimeError("reference binary failed: " + lg[-1500:])
        ph = [bm_ref.parse_log(lg) for lg in logs]
        for x in ph:
            for key in ("map_s", "kmer_index_s", "find_offset_s", "output_s", "load_s"):
                x.setdefault(key, 0.0)
            x.setdefault("locations", 0)
        per = [x["map_s"] + x["kmer_index_s"] + x["find_offset_s"] + (x["output_s"] if align else 0.0) for x in ph]
        t = max(per)
        if keep is not None:
            for pi, (seq, qual, off) in enumerate(inputs):
                keep.append((seq, qual, off, open(os.path.join(work, f"p{pi}", "out.sam")).read()))
        return {"reads": procs * n_per_proc, "seconds": t, "procs": procs, "reads_per_s": procs * n_per_proc / t,
                "map_s": max(x["map_s"] for x in ph), "kmer_index_s": max(x["kmer_index_s"] for x in ph),
                "find_offset_s": max(x["find_offset_s"] for x in ph), "load_s": max(x["load_s"] for x in ph),
                "output_s": max(x["output_s"] for x in ph), "align": align,
                "wall_s": wall, "span": span, "n_locs": sum(x["locations"] for x in ph)}
    finally:
        shutil.rmtree(work, ignore_errors=True)


def ref_procs():
    ncores = os.cpu_count() or 1
    try:
        import psutil
        by_mem = int(psutil.virtual_memory().available // (6 << 30))      # ~4 GB peak per reference process at C2
    except Exception:
        by_mem = 8
    return max(1, min(ncores, 16, by_mem))


def ref_sample_text(c):
    return (f"{c['procs']} single-threaded reference processes (oracle/_ref, the reference's own main.cpp) x "
            f"{c['reads'] // c['procs']} reads, each drawn from its own {c['span']}-base span (same reads per bucket as "
            f"the full batch), full index; slowest process: bucket mapping {c['map_s']:.2f}s + per-bucket k-mer index "
            f"{c['kmer_index_s']:.2f}s + offset vote {c['find_offset_s']:.2f}s"
            + (f" + alignment verification and output {c['output_s']:.2f}s" if c.get("align") else "")
            + f" (its own [BENCHMARK] timers; index load {c['load_s']:.0f}s and FASTA parse not counted)")


def sam_parity(ctx, bm, P, L, kept, align):
    """the SAM records the reference processes wrote for their sample reads against the records the library returns for
    the same reads: QNAME, FLAG, RNAME, POS, MAPQ (and CIGAR for the align flavour) of every record, in order"""
    n_reads = n_rec = 0
    for seq, qual, off, sam in kept:
        want = []
        for line in sam.splitlines():
            if not line or line[0] == "@":
                continue
            f = line.split("\t", 6)
            want.append((int(f[0]), int(f[1]), f[2], int(f[3]), int(f[4]), f[5]))
        n = len(off) - 1
        if not align:
            loff, locs, _ = ctx.map_locate(P, seq, qual, off, True)
            rid = np.repeat(np.arange(n), np.diff(loff).astype(np.int64))
            got = list(zip(rid.tolist(), np.where(locs["is_original"] == 1, 0, 16).tolist(), ["chr1"] * len(locs),
                           (locs["bucket"].astype(np.int64) * L + locs["offset"] + 1).tolist(),
                           np.minimum(60, 6 * locs["votes"].astype(np.int64)).tolist(), ["*"] * len(locs)))
        else:
            ctx.reads_upload(P, seq, qual, off)
            ctx.map_device(P)
            ctx.locate_device(P, False)
            loff, locs = ctx.locate_download()
            ctx.align_device(P)
            aln, cig = ctx.align_download(len(locs))
            rid = np.repeat(np.arange(n), np.diff(loff).astype(np.int64))
            got = []
            for j in np.nonzero(aln["kept"])[0].tolist():
                ops = cig[aln["cigar_off"][j]:aln["cigar_off"][j] + aln["cigar_len"][j]]
                text = "".join("%d%s" % (int(o) >> 4, "MID"[int(o) & 15]) for o in ops) or "*"
                got.append((int(rid[j]), 0 if locs["is_original"][j] else 16, "chr1",
                            int(aln["begin"][j]) + int(locs["bucket"][j]) * L + int(locs["offset"][j]) + 1,
                            int(aln["map_qual"][j]) & 0xFF, text))
        if got != want:
            bad = next((i for i, (x, y) in enumerate(zip(got, want)) if x != y), min(len(got), len(want)))
            return {"reads": n_reads + n, "records": n_rec, "sam_fields_equal": False,
                    "first_difference": {"record": bad, "got": str(got[bad:bad + 1]), "want": str(want[bad:bad + 1]),
                                         "n_got": len(got), "n_want": len(want)}}
        n_reads += n
        n_rec += len(want)
    return {"reads": n_reads, "records": n_rec, "sam_fields_equal": True,
            "fields": "QNAME FLAG RNAME POS MAPQ CIGAR of every record, in file order",
            "against": "the SAM files written by the reference processes of the cpu_baseline leg (oracle/_ref)"}


def main():
    a = parse()
    cfg = CONFIGS[a.config]
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    align = bool(cfg.get("align"))
    reads_total = a.reads or cfg["reads"]
    n_reads_cfg = reads_total // world if (cfg.get("strong") and not a.reads) else reads_total
    workload = f"{a.config}: {cfg['desc']}, -f {a.frac:g}; {n_reads_cfg} reads per GPU per step"
    metric, unit = "reads/sec mapped", "reads/s"

    if a.impl == "reference":
        if rank != 0:
            return 0
        run_reference(a, cfg, workload, metric, unit)
        return 0

    import torch
    import torch.distributed as dist
    import bucketmap_b200 as bm

    # libraries (NCCL) print banners on stdout: keep fd 1 for the one JSON line
    real_stdout = os.dup(1)
    os.dup2(2, 1)
    torch.cuda.set_device(local)
    all_cpus = os.sched_getaffinity(0)
    bind_to_gpu_numa_node(local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))   # rendezvous, barriers, max over ranks
    ctx = bm.Context(local)
    ctx.index_kmer_table(a.ktable)
    ctx.index_fine_table(a.ftable)
    ctx.locator_index(a.lidx)
    ctx.pipeline_lanes(a.lanes)
    q, L = cfg["q"], cfg["L"]
    P = bm.default_params(ctx.lib, L, q=q)

    # ---- reference + index: every rank uploads the (seeded) genome; rank 0 builds the matrix; the library replicates it
    t0 = time.perf_counter()
    packed, NB, b_start, b_len = make_workload(cfg)
    k2i, n_rows = fracminhash(q, FMH_X, FMH_Y, a.frac)
    ctx.ref_upload(packed, cfg["n_bases"], b_start, b_len, NB)
    ctx.kernel_timing(True)
    sharded_ms = sharded_ok = None
    if world > 1:
        uid = [ctx.comm_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(uid, src=0)
        ctx.comm_init(uid[0], world, rank)               # ncclCommInitRank inside the library (dlopen'ed NCCL)
        # K1 sharded by buckets + ncclAllReduce (bm_index_build_sharded), timed, then checked against the one-rank build that
        # the broadcast below delivers: the per-row zero counts of both matrices must agree on every rank
        torch.cuda.synchronize()
        dist.barrier()
        ctx.index_build_sharded(k2i, q, n_rows)          # the first collective also brings NCCL's channels up: not timed
        torch.cuda.synchronize()
        dist.barrier()
        tb = time.perf_counter()
        ctx.index_build_sharded(k2i, q, n_rows)
        dist.barrier()
        sharded_ms = (time.perf_counter() - tb) * 1e3
        zeros_sharded = ctx.index_zeros(0.5)[0]
        ctx.kernel_stats_reset()
    if rank == 0:
        ctx.index_build(k2i, q, n_rows, download=False)
        ctx.sync()
    index_ms = ctx.kernel_stats()["index_build"][0]
    bcast_ms = None
    if world > 1:
        torch.cuda.synchronize()
        dist.barrier()
        tb = time.perf_counter()
        ctx.bcast_index(0)                               # ncclBroadcast: geometry, kmer_to_index, bit matrix
        dist.barrier()
        bcast_ms = (time.perf_counter() - tb) * 1e3
        ctx.n_rows = n_rows
        okt = torch.tensor([1 if np.array_equal(zeros_sharded, ctx.index_zeros(0.5)[0]) else 0], device="cuda")
        dist.all_reduce(okt, op=dist.ReduceOp.MIN)
        sharded_ok = bool(okt.item())
    W = (NB + 7) >> 3

    # ---- reads of this rank (its contiguous shard of the job's reads), in both host forms
    from bucketmap_b200 import datagen
    seq, qual, off, truth_start, truth_rev = datagen.simulate_reads(packed, cfg["n_bases"], n_reads_cfg, cfg["read_len"], cfg["sub"],
                                                                    cfg["indel"], cfg["indel"], cfg["rseed"] + rank)
    n_reads = len(off) - 1
    planes, qk, pl_off, n_low = bm.pack_reads(ctx.lib, seq, qual, off, P.k, P.avg_base_quality)
    setup_s = time.perf_counter() - t0

    def barrier():
        ctx.sync()
        if world > 1:
            ctx.gather_sync()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()

    gather_verified = None
    if world > 1:
        hs = [None] * world
        dist.all_gather_object(hs, ctx.gather_open(rank, world, 0, n_reads, 4 * n_reads + 4096))
        ctx.gather_connect(hs)

    # ---- device-resident steps ------------------------------------------------------------------------------------
    ctx.reads_upload_packed(P, planes, qk if n_low else None, pl_off, off)

    def step_resident():
        ctx.map_device(P)
        ctx.locate_device(P, not align)
        if align:
            ctx.align_device(P)
        if world > 1:
            ctx.gather_results()                 # records -> rank 0's device memory, behind the next step's mapper

    for _ in range(a.warmup):
        step_resident()
    barrier()
    if world > 1:
        # rank 0 re-maps the shard of the last rank locally and compares it with what the gather delivered
        if rank == 0:
            goff, glocs = ctx.gather_download()
            r = world - 1
            s2, q2, o2 = simulate(cfg, packed, n_reads_cfg, cfg["rseed"] + r)
            ctx.reads_upload(P, s2, q2, o2)
            ctx.map_device(P)
            ctx.locate_device(P, not align)
            eoff, elocs = ctx.locate_download()
            lo = r * n_reads
            ok = (len(goff) == world * n_reads + 1 and
                  np.array_equal(goff[lo:lo + n_reads + 1].astype(np.int64) - int(goff[lo]), eoff.astype(np.int64)) and
                  np.array_equal(glocs[int(goff[lo]):int(goff[lo + n_reads])], elocs))
            ctx.locate_device(P, not align)      # own shard back in the result buffers
            ctx.reads_upload_packed(P, planes, qk if n_low else None, pl_off, off)
            ctx.map_device(P); ctx.locate_device(P, not align)
            gather_verified = bool(ok)
        barrier()
    ktable_ms = ctx.kernel_stats()["kmer_table"][0]     # one-off, inside the first (untimed) mapper call
    lidx_ms = ctx.kernel_stats()["locator_index"][0]    # one-off, inside the first (untimed) locator call
    ktable_k, ktable_bytes = ctx.index_kmer_table_info()
    ftable_k, ftable_bytes = ctx.index_fine_table_info()
    ktable_group, ktable_density = ctx.index_kmer_table_choice()
    ctx.kernel_stats_reset()
    barrier()
    clocks = ClockSampler(local)
    clocks.start()
    ctx.timer_start()
    for _ in range(a.steps):
        step_resident()
    if world > 1:
        ctx.gather_join()                # the library stream waits for the gather stream: the events cover its tail
    ms = ctx.timer_stop()                # CUDA events on the library's stream: every kernel of the path
    barrier()
    clk = clocks.stop()
    kst = ctx.kernel_stats()
    _, _, mstats = ctx.map_download()
    res_off, res_locs = ctx.locate_download()
    n_aln = len(res_locs) if align else 0
    acc = accuracy(res_off, res_locs, L, truth_start, truth_rev) if rank == 0 else None
    tms = torch.tensor([ms], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(tms, op=dist.ReduceOp.MAX)
    ms_max = float(tms.item())
    value = world * n_reads * a.steps / (ms_max * 1e-3)

    # ---- end to end: pinned host buffers -> bm_map_locate_packed -> host records ---------------------------------------
    e2e_steps = a.e2e_steps or a.steps
    hpl = ctx.pinned(len(planes), np.uint64); hpl[:] = planes
    hqk = None
    if n_low:
        hqk = ctx.pinned(len(qk), np.uint32); hqk[:] = qk
    hplo = ctx.pinned(len(pl_off), np.uint32); hplo[:] = pl_off
    hoff = ctx.pinned(len(off), np.uint64); hoff[:] = off
    ooff = ctx.pinned(world * n_reads + 1 if rank == 0 else n_reads + 1, np.uint32)
    olocs = ctx.pinned(max(4 * n_reads * (world if rank == 0 else 1), 1024), bm.LOC_DTYPE)   # merged best locations: ~1 per read
    h2d_packed = int(pl_off[-1]) * 8 + (int(pl_off[-1]) * 4 if n_low else 0) + hplo.nbytes + hoff.nbytes
    d2h = 0
    e2e = e2e_ascii = e2e_one = e2e_gathered = None
    if not align:
        import threading

        gathered = [False]                       # N > 1: True = the records of all ranks go to rank 0's HOST through the gather

        def step_e2e(c, oo, ol):
            if not gathered[0]:
                # every rank: its shard from its host buffers, its records into its own host arrays (what a sharded job needs:
                # each process writes the SAM records of its reads, as the reference arm's processes do)
                n_locs, _ = c.map_locate_packed_into(P, hpl, hqk, hplo, hoff, oo[:n_reads + 1], ol[:4 * n_reads], True)
                return (n_reads + 1) * 4 + n_locs * bm.LOC_DTYPE.itemsize
            c.map_locate_packed_into(P, hpl, hqk, hplo, hoff, None, None, True)
            if rank == 0:
                nr, nl = c.gather_results(oo, ol)
                return (nr + 1) * 4 + nl * bm.LOC_DTYPE.itemsize
            c.gather_results()
            return 0

        def run_e2e(lanes):
            """lanes: [(context, out offsets, out records)], one host thread each; every step is one call with its own H2D + D2H"""
            def sync_all():
                for c, _, _ in lanes:
                    c.sync()
                    if gathered[0]:
                        c.gather_sync()
                torch.cuda.synchronize()
                if world > 1:
                    dist.barrier()
            per_lane = (e2e_steps + len(lanes) - 1) // len(lanes)
            got = [0] * len(lanes)
            errs = []

            def drive(i, steps):
                try:
                    for _ in range(steps):
                        got[i] = step_e2e(*lanes[i])
                except Exception as e:               # noqa: BLE001 -- re-raised by the main thread
                    errs.append(e)
            for i in range(len(lanes)):
                drive(i, max(1, min(a.warmup, 3)))
            sync_all()
            te = time.perf_counter()
            if len(lanes) == 1:
                drive(0, per_lane)
            else:
                th = [threading.Thread(target=drive, args=(i, per_lane)) for i in range(len(lanes))]
                for t in th:
                    t.start()
                for t in th:
                    t.join()
            sync_all()
            secs = time.perf_counter() - te
            if errs:
                raise errs[0]
            t2 = torch.tensor([secs], dtype=torch.float64, device="cuda")
            if world > 1:
                dist.all_reduce(t2, op=dist.ReduceOp.MAX)
            return world * n_reads * per_lane * len(lanes) / float(t2.item()), per_lane * len(lanes), got[0]

        api = ("bm_map_locate_packed (pinned host buffers in the FASTQ parser's packed form: 2-bit base planes" +
               (", k-mer quality plane" if n_low else ", no quality plane: no read has a base below -b") +
               "; records into the rank's own pinned host arrays; H2D chunks pipelined against the mapper inside the call)")

        def line(v, steps, d2h_b, what, **more):
            out = {"value": v, "unit": unit, "h2d_bytes_per_step": int(h2d_packed) * world,
                   "d2h_bytes_per_step": int(d2h_b) * (1 if gathered[0] else world), "steps": steps, "api": api + what}
            out.update(more)
            return out
        v1, s1, d2h = run_e2e([(ctx, ooff, olocs)])
        e2e_one = line(v1, s1, d2h, "; one call at a time")
        e2e = e2e_one
        e2e_gathered = None
        if a.e2e_lanes != 1:
            # two batches in flight per GPU: the context and a lane (bm_ctx_create_lane: views the index, owns streams and
            # batch buffers), one host thread each
            lane = ctx.create_lane(P)
            ooff2 = ctx.pinned(n_reads + 1, np.uint32)
            olocs2 = ctx.pinned(4 * n_reads, bm.LOC_DTYPE)
            v2, s2, d2h2 = run_e2e([(ctx, ooff, olocs), (lane, ooff2, olocs2)])
            same = bool(np.array_equal(ooff[:n_reads + 1], ooff2) and
                        np.array_equal(olocs[:int(ooff[n_reads])], olocs2[:int(ooff2[-1])]))
            e2e = line(v2, s2, d2h2, "; two calls in flight per GPU: the context and a lane (bm_ctx_create_lane), one host thread each",
                       lanes_give_the_same_records=same)
            lane.close()
        if world > 1:
            # the same with all records delivered to rank 0's host (bm_gather_results + bm_gather_host_async): what one writer of
            # one SAM file in read order needs; bound by rank 0's link, which carries every rank's records
            gathered[0] = True
            if rank == 0:
                ctx.gather_host_async(True)
            vg, sg, d2hg = run_e2e([(ctx, ooff, olocs)])
            e2e_gathered = line(vg, sg, d2hg, "; + bm_gather_results: every rank's records into rank 0's host memory; one call at a time")
            gathered[0] = False
        if world == 1:
            hseq = ctx.pinned(len(seq), np.uint8); hseq[:] = seq
            hqual = ctx.pinned(len(qual), np.uint8); hqual[:] = qual
            for _ in range(2):
                ctx.map_locate_into(P, hseq, hqual, hoff, ooff, olocs, True)
            barrier()
            te = time.perf_counter()
            na = max(1, min(e2e_steps, 5))
            for _ in range(na):
                ctx.map_locate_into(P, hseq, hqual, hoff, ooff, olocs, True)
            barrier()
            e2e_ascii = {"value": n_reads * na / (time.perf_counter() - te), "unit": unit,
                         "h2d_bytes_per_step": int(hseq.nbytes + hqual.nbytes + hoff.nbytes), "steps": na,
                         "api": "bm_map_locate (ASCII bases + qualities as the FASTQ holds them; packed on the device)"}
    else:
        # bucketmap_align: upload + map + locate + verify + download of the alignment records into pinned host arrays; two
        # batches in flight (the context and a lane, one host thread each) unless --e2e-lanes 1
        import threading
        cap_l = 2 * n_reads + 4096

        def make_out(c):
            return (c.pinned(n_reads + 1, np.uint32), c.pinned(cap_l, bm.LOC_DTYPE), c.pinned(cap_l, bm.ALN_DTYPE),
                    c.pinned(8 * cap_l, np.uint32))

        def step_e2e(c, o):
            c.reads_upload_packed(P, hpl, hqk, hplo, hoff)
            c.map_device(P); c.locate_device(P, False); c.align_device(P)
            nl = c.locate_download_into(o[0], o[1])
            used = c.align_download_into(o[2][:nl], o[3])
            return o[0].nbytes + nl * (bm.LOC_DTYPE.itemsize + bm.ALN_DTYPE.itemsize) + used * 4

        lanes = [(ctx, make_out(ctx))]
        if a.e2e_lanes != 1:
            ln = ctx.create_lane(P)
            lanes.append((ln, make_out(ctx)))
        got = [0] * len(lanes)
        errs = []

        def drive(i, steps):
            try:
                for _ in range(steps):
                    got[i] = step_e2e(*lanes[i])
            except Exception as e:               # noqa: BLE001 -- re-raised by the main thread
                errs.append(e)
        for i in range(len(lanes)):
            drive(i, 1)
        for c, _ in lanes:
            c.sync()
        barrier()
        per_lane = (e2e_steps + len(lanes) - 1) // len(lanes)
        te = time.perf_counter()
        th = [threading.Thread(target=drive, args=(i, per_lane)) for i in range(len(lanes))]
        for t in th:
            t.start()
        for t in th:
            t.join()
        for c, _ in lanes:
            c.sync()
        barrier()
        secs = time.perf_counter() - te
        if errs:
            raise errs[0]
        same = None
        if len(lanes) == 2:
            n0 = int(lanes[0][1][0][-1])
            same = bool(all(np.array_equal(x[:n0] if i in (1, 2) else x, y[:n0] if i in (1, 2) else y)
                            for i, (x, y) in enumerate(zip(lanes[0][1][:3], lanes[1][1][:3]))))
            lanes[1][0].close()
        e2e = {"value": n_reads * per_lane * len(lanes) / secs, "unit": unit, "h2d_bytes_per_step": int(h2d_packed),
               "d2h_bytes_per_step": int(got[0]), "steps": per_lane * len(lanes), "lanes_give_the_same_records": same,
               "api": "bm_reads_upload_packed + bm_map_device + bm_locate_device + bm_align_device + bm_locate_download + "
                      "bm_align_download into pinned arrays" + ("; two batches in flight: the context and a lane" if len(lanes) == 2 else "")}

    # ---- roofline of the dominant kernel -------------------------------------------------------------------------------
    peak, peak_src = measured_peak()
    vote_ms, vote_n = kst["map_vote"]
    alt = None
    if ktable_k and rank == 0 and world == 1 and not align and a.config in ("c2", "c2r", "c2x"):
        # the full-width q-gram gather (k_map_vote) on the same batch, outside the timed region: the mapper when the
        # tables are off / do not fit, and the kernel the HBM roofline target of BASELINE.json was written for
        ctx.index_kmer_table(0)
        ctx.reads_upload_packed(P, planes, qk if n_low else None, pl_off, off)
        ctx.map_device(P)                                # warm-up (frees the tables)
        ctx.kernel_stats_reset()
        for _ in range(3):
            ctx.map_device(P)
        ctx.sync()
        k2 = ctx.kernel_stats()
        _, _, ms2 = ctx.map_download()
        t_ms = k2["map_vote"][0] / max(1, k2["map_vote"][1])
        alt = {"kernel": "k_map_vote", "bound": "hbm", "achieved": int(ms2[4]) / (t_ms * 1e-3) / 1e9,
               "unit": "GB/s", "alg_bytes_per_launch": int(ms2[4]), "ms_per_launch": t_ms,
               "map_sample_ms_per_launch": k2["map_sample"][0] / max(1, k2["map_sample"][1]),
               "alg_bytes_definition": "16-B q-gram row loads issued by k_map_vote (counted in-kernel) until all 4096 "
                                       "buckets of a 512-B column tile have missed F samples",
               "survey_8d_bytes_per_launch": int(ms2[3]) * W + n_reads * 2 * cfg["read_len"]}
        alt["peak"] = peak; alt["frac"] = alt["achieved"] / peak
        ctx.index_kmer_table(a.ktable)
    rows_gathered = int(mstats[3])          # rows the reference's cascade ANDs (R of SURVEY 8d)
    alg_bytes = int(mstats[4])              # bytes the mapper's loads asked for, counted in the kernel
    launches = int(sum(v[1] for v in kst.values()))
    if align:
        align_ms = kst["align"][0] / max(1, kst["align"][1]) * (kst["align"][1] / a.steps)
        width = cfg["read_len"] + 1 + int(P.indel_rate * cfg["read_len"])
        cells = n_aln * cfg["read_len"] * width
        roof = {"kernel": "k_align", "bound": "integer ALU (no HBM / tensor roofline applies; SURVEY 8d: report GCUPS)",
                "achieved": cells / (align_ms * 1e-3) / 1e9, "peak": None, "unit": "GCUPS", "frac": None, "traffic": None,
                "alignments_per_step": n_aln, "cells_per_alignment": cfg["read_len"] * width, "ms_per_step": align_ms,
                "alg_bytes_definition": "cell updates = alignments x read length x text window columns (len + 1 + "
                                        "(size_t)(n * len), bucket_locator.h:550)"}
    else:
        kname = "k_map_fused" if ktable_k else "k_map_vote"
        achieved = alg_bytes / (vote_ms / vote_n * 1e-3) / 1e9 if vote_n else None
        traffic, traffic_src = ncu_traffic(a.config, a.frac, n_reads, kname)
        t_s = vote_ms / vote_n * 1e-3 if vote_n else None
        # the same launch in DRAM lines, from THIS run's counters (no stored capture involved): a coarse row costs its whole
        # 128-B lines, a recount word the 64-B sector its load hints at, the planes their bytes
        model = None
        if ktable_k and ktable_group:
            S = int(P.mapper_samples)
            row_lines = (((NB + ktable_group - 1) // ktable_group + 7) // 8 + 127) // 128
            n_lists = 2 * n_reads                                   # one segment per read in every bench configuration
            recount_words = int(mstats[6]) * S * (1 if ftable_k else (P.k - q + 1))
            plane_bytes = alg_bytes - n_lists * S * (((NB + ktable_group - 1) // ktable_group + 31) // 32) * 4 - recount_words * 4
            model = n_lists * S * row_lines * 128 + recount_words * 64 + max(0, plane_bytes)
        roof = {"kernel": kname, "bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s",
                "frac": (achieved / peak) if achieved else None, "traffic": traffic, "traffic_source": traffic_src,
                "traffic_gbs": (traffic / t_s / 1e9) if (traffic and t_s) else None,
                "traffic_frac": (traffic / t_s / 1e9 / peak) if (traffic and t_s) else None,
                "traffic_model": model, "traffic_model_frac": (model / t_s / 1e9 / peak) if (model and t_s) else None,
                "traffic_model_definition": "DRAM lines the launch must move, from this run's in-kernel counters: 128 B per line of a "
                                            "coarse row, 64 B per recount word, the read planes; `traffic` is the ncu measurement of "
                                            "the same launch, valid while its source hash matches",
                "peak_source": peak_src, "alg_bytes_per_launch": alg_bytes, "ms_per_launch": vote_ms / vote_n if vote_n else None,
                "rows_gathered_per_read": rows_gathered / max(1, n_reads),
                "survey_8d_bytes_per_launch": rows_gathered * W + n_reads * 2 * cfg["read_len"],
                "survey_8d_note": "SURVEY 8(d)'s B_map = R*W + read bytes is the q-gram cascade the reference runs (and "
                                  "k_map_vote implements, see roofline_qgram_path: its early exit already needs fewer bytes); "
                                  "k_map_fused replaces that gather by one precomputed coarse k-mer row per sample plus an "
                                  "exact recount of the passing groups and is NOT measured against those bytes: `frac` is "
                                  "on the bytes its own loads ask for, `traffic_frac` on what crossed the DRAM pins. The "
                                  ">=60% HBM target of BASELINE.json is met by traffic_frac and by k_map_vote's frac, not by "
                                  "k_map_fused's algorithmic frac",
                "alg_bytes_definition": ("bytes k_map_fused's loads ask for (counted in-kernel): the segment's bit planes (8 B "
                                         "per 32 bases, +4 B with a quality plane) + S coarse k-mer rows per strand (row bytes "
                                         "in use, not the 256-B stride) + 4 B per table word read for the exact recount of "
                                         "passing groups") if ktable_k else
                                        "16-B q-gram row loads issued by k_map_vote (counted in-kernel)"}
    out = {
        "metric": metric, "value": value, "unit": unit, "n_gpus": world, "steps": a.steps, "warmup": a.warmup,
        "ms_per_step": ms_max / a.steps, "higher_is_better": True, "scaling": "strong" if (cfg.get("strong") and not a.reads) else "weak",
        "vs_baseline": None, "dtype": "u32 bitwise" if not align else "u64 bitwise", "data": "synthetic",
        "config": {"workload": workload, "genome_bases": cfg["n_bases"], "bucket_len": L, "NB": NB,
                   "row_bytes": W, "matrix_bytes": int(n_rows) * W, "kmer_table_bytes": int(ktable_bytes),
                   "fine_table_bytes": int(ftable_bytes),
                   "kmer_table": ("4^%d coarse k-mer rows (AND of each k-mer's q-gram rows, OR-reduced over groups of %d buckets: "
                                  "the cost model's choice at a measured row density of %.4f)%s, derived on the device from the "
                                  "matrix in %.1f ms; the mapper gathers 1 coarse row per sample and recounts passing groups exactly" % (
                                      ktable_k, ktable_group, ktable_density,
                                      " + the same rows at full bucket resolution for the recount" if ftable_k else "", ktable_ms))
                                 if ktable_k else ("none: the mapper gathers the k-q+1 q-gram rows per sample (k_map_sample + k_map_vote)"
                                                   + (("; the cost model found the gather cheaper at a k-mer row density of %.3f" % ktable_density)
                                                      if (a.ktable == 1 and ktable_density) else "")),
                   "kmer_table_buckets_per_bit": int(ktable_group), "kmer_row_density": ktable_density,
                   "reads_per_gpu_per_step": n_reads, "read_len": cfg["read_len"],
                   "flags": "-k %d -l 12 -s 15 -e 0.4 -p 10 -f %g" % (q, a.frac),
                   "parallelism": f"reads sharded x{world}, index replicated",
                   "timing": "CUDA events on the library stream" + ("; the stream joins the gather stream before the closing event; max over ranks" if world > 1 else ""),
                   "l2": "inputs larger than L2 (coarse k-mer table %d MB, bit matrix %d MB vs 126 MB L2)" % (ktable_bytes >> 20, n_rows * W >> 20),
                   "resident_input": "packed read batch (2-bit base planes%s), as bm_reads_upload_packed / the parser leave it" % (
                       " + k-mer quality plane" if n_low else "; no quality plane: every base >= -b"),
                   "locator": "genome k-mer position index (K3i)" if a.lidx else "bucket text scan (K3)",
                   "step": ("bm_map_device + bm_locate_device(filter_best=1)" if not align else
                            "bm_map_device + bm_locate_device(filter_best=0) + bm_align_device") +
                           (" + bm_gather_results (records into rank 0's device memory)" if world > 1 else "")},
        "e2e": e2e, "e2e_one_call_at_a_time": e2e_one, "e2e_gathered_to_rank0": e2e_gathered, "e2e_ascii": e2e_ascii,
        "gpu_launches": launches,
        "clocks": clk,
        "roofline": roof,
        "roofline_qgram_path": alt,
        "kernels_ms_per_step": {k: v[0] / a.steps for k, v in kst.items() if v[1]},
        "index_build_ms": index_ms, "kmer_table_build_ms": ktable_ms, "locator_index_build_ms": lidx_ms,
        "index_bcast_ms": bcast_ms, "index_build_sharded_ms": sharded_ms, "index_build_sharded_verified": sharded_ok,
        "setup_s": setup_s,
        "mapped": {"reads_with_candidate": int(mstats[0]), "cand_fwd": int(mstats[1]), "cand_rev": int(mstats[2]),
                   "lists_handed_to_full_width_gather": int(mstats[5]), "coarse_groups_recounted_per_read": int(mstats[6]) / max(1, n_reads),
                   "locator_candidates_to_bucket_scan": int(mstats[7])},
    }
    out["accuracy"] = acc
    if world > 1:
        out["gather_verified"] = gather_verified
        out["gather"] = "bm_gather_results: peer-memory stores over NVLink into rank 0's slots + cuStreamWaitValue32 + pack kernel"
    if rank == 0 and world == 1 and not a.no_cpu:
        os.sched_setaffinity(0, all_cpus)            # the CPU baseline may use every host core
        rows = ctx.index_download()
        ncores = os.cpu_count() or 1
        sys.path.insert(0, os.path.join(ROOT, "oracle"))
        import bm_ref
        if bm_ref.available(NB, L, align):
            kept = []
            c = reference_leg(cfg, a.frac, packed, NB, b_start, k2i, rows, a.cpu_sample or (12_000 if not align else 6_000),
                              n_reads_cfg, ref_procs(), keep=kept)
            out["cpu_baseline"] = {"value": c["reads_per_s"], "unit": unit, "cores": c["procs"], "kind": "reference",
                                   "sample": ref_sample_text(c)}
            out["parity"] = sam_parity(ctx, bm, P, L, kept, align)
        else:
            n_sample = a.cpu_sample or max(2000, min(50_000, 600 * ncores))
            c = cpu_leg(cfg, a.frac, packed, NB, b_start, b_len, k2i, rows, n_sample, n_reads_cfg, ncores)
            out["cpu_baseline"] = {"value": c["reads_per_s"], "unit": unit, "cores": c["threads"], "kind": "port",
                                   "sample": f"{c['reads']} reads drawn from the first {c['span']} bases (same reads per "
                                             f"bucket as the full batch), full index; map {c['map_s']:.2f}s + locate/merge "
                                             f"{c['locate_s']:.2f}s"}
    if rank == 0:
        os.write(real_stdout, (json.dumps(out) + "\n").encode())
    if world > 1:
        ctx.gather_close()
        dist.barrier()
        dist.destroy_process_group()
    ctx.close()
    return 0


def fracminhash(q, x, y, frac):
    """tools/hash_function_generator.h:95-117 + indexer/bucket_indexer.h:147-159 (host logic of the product)"""
    from bucketmap_b200 import hostlogic
    return hostlogic.fracminhash(q, x, y, frac)


def run_reference(a, cfg, workload, metric, unit):
    """reference arm: the reference's own CPU implementation of the path on the host cores.  oracle/_ref (the
    reference's main.cpp built from /root/reference against oracle/shim) when the binary for this bucket count was
    prebuilt, else the oracle port with all host threads."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import bm_oracle
    import bm_ref
    bm_oracle.build()
    ncores = os.cpu_count() or 1
    o = bm_oracle.Oracle(ncores)
    q, L, align = cfg["q"], cfg["L"], bool(cfg.get("align"))
    full_reads = a.reads or cfg["reads"]
    packed, NB, b_start, b_len = make_workload(cfg)
    k2i, n_rows = o.fracminhash(q, FMH_X, FMH_Y, a.frac)
    rows = o.index_build(packed, b_start, b_len, NB, q, k2i, n_rows)   # setup only: == `bucketmap -x` (tests/golden)
    steps = max(1, a.steps)
    if bm_ref.available(NB, L, align):
        n_per = a.cpu_sample or max(2000, min(12_000, 36_000 // steps))
        c = reference_leg(cfg, a.frac, packed, NB, b_start, k2i, rows, n_per * steps, full_reads, ref_procs())
        v = c["reads_per_s"]
        print(json.dumps({
            "impl": "reference", "metric": metric, "value": v, "unit": unit, "n_gpus": a.gpus, "steps": steps,
            "warmup": a.warmup, "ms_per_step": c["seconds"] / steps * 1e3, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "u64 bitwise (std::bitset)", "data": "synthetic",
            "config": {"workload": workload, "NB": NB,
                       "note": "unmodified reference sources (oracle/_ref); a step = %d reads per process; one launch "
                               "covers all steps, warm-up reads are not separable inside the reference process" % n_per},
            "cpu_baseline": {"value": v, "unit": unit, "cores": c["procs"], "kind": "reference", "sample": ref_sample_text(c)},
            "e2e": {"value": v, "unit": unit, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        }), flush=True)
        return
    P = o.params(L, q=q)
    n_sample = a.cpu_sample or max(2000, min(50_000, 600 * ncores))
    span = min(max(int(cfg["n_bases"] * n_sample / full_reads), 4 * L), cfg["n_bases"])
    seq, qual, off = simulate(cfg, packed, n_sample, cfg["rseed"] + 1000, n_span=span)
    segs = o.segments(P, off)
    ix = o.index_open(P, NB, k2i, rows)

    def step():
        coff, cand, _ = o.map_segments_h(ix, P, seq, qual, off, segs)
        loff, locs = o.locate(P, packed, b_start, b_len, seq, qual, off, segs, coff, cand)
        if align:
            return o.align(P, packed, b_start, b_len, seq, off, loff, locs)
        return o.filter_best(P, off, loff, locs)

    for _ in range(min(a.warmup, 1)):
        step()
    steps = max(1, min(a.steps, 3))
    t0 = time.perf_counter()
    for _ in range(steps):
        step()
    dt = time.perf_counter() - t0
    v = n_sample * steps / dt
    sample = (f"{n_sample} reads per step drawn from the first {span} bases (same reads per bucket as the full "
              f"batch), full index, {steps} timed steps")
    print(json.dumps({
        "impl": "reference", "metric": metric, "value": v, "unit": unit, "n_gpus": a.gpus, "steps": steps,
        "warmup": min(a.warmup, 1), "ms_per_step": dt / steps * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u64 bitwise", "data": "synthetic",
        "config": {"workload": workload, "NB": NB, "note": "CPU restatement of the reference (oracle/), all host threads"},
        "cpu_baseline": {"value": v, "unit": unit, "cores": o.threads, "kind": "port", "sample": sample},
        "e2e": {"value": v, "unit": unit, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }), flush=True)


if __name__ == "__main__":
    sys.exit(main())
